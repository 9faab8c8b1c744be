/*
 * oracle/synth.c -- TEST INFRASTRUCTURE (see pcs_oracle.h).  Synthetic organized RGB-D scan of SURVEY 8(d).
 *
 * The reference ships no scan data (its demo bags are missing, .MISSING_LARGE_BLOBS:1-3), so every measurement
 * input is synthetic.  The generator is written with explicitly ordered double arithmetic and a polynomial
 * sine (no libm in the per-pixel path) so that the CUDA generator in point_cloud_segmentation_b200/csrc/synth.cu
 * produces bit-identical frames; build with -ffp-contract=off.
 *
 * Scene: bumpy plane z_w = 0.02 sin(8 x_w) cos(6 y_w) seen by a pinhole camera (fx = fy = 0.9 W) 0.8 m above it,
 * looking straight down, moving +4 mm in x per frame with a small yaw; 5 % NaN holes; uniform-random background
 * colours with painted stripes whose colour falls inside the reference's HSV thresholds
 * (pcs_detection/include/pcs_detection/hsv_thresholding.h:47-52).
 */
#include "pcs_oracle.h"

#include <math.h>
#include <string.h>

uint64_t orc_splitmix64(uint64_t x)
{
  x += 0x9E3779B97F4A7C15ULL;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ULL;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBULL;
  return x ^ (x >> 31);
}

double orc_psin(double x)
{
  /* reduce to r in [-pi, pi], then an odd Taylor polynomial to x^15 (abs error < 1e-6: plenty for a test scene) */
  const double inv_2pi = 0.15915494309189535, two_pi = 6.283185307179586;
  const double k = floor(x * inv_2pi + 0.5);
  const double r = x - k * two_pi;
  const double r2 = r * r;
  double p = -1.0 / 1307674368000.0;
  p = p * r2 + 1.0 / 6227020800.0;
  p = p * r2 + -1.0 / 39916800.0;
  p = p * r2 + 1.0 / 362880.0;
  p = p * r2 + -1.0 / 5040.0;
  p = p * r2 + 1.0 / 120.0;
  p = p * r2 + -1.0 / 6.0;
  p = p * r2 + 1.0;
  return p * r;
}

double orc_pcos(double x) { return orc_psin(x + 1.5707963267948966); }

void orc_synth_pose(int frame, float pose[16])
{
  /* camera -> world, row-major: R = Rz(yaw) * diag(1,-1,-1), t = (0.004 f, 0, 0.8).  Host-only (libm). */
  const double yaw = (0.2 * 3.141592653589793 / 180.0) * sin((double)frame / 50.0);
  const float c = (float)cos(yaw), s = (float)sin(yaw);
  const float m[16] = { c, s, 0.f, (float)(0.004 * (double)frame), s, -c, 0.f, 0.f, 0.f, 0.f, -1.f, 0.8f, 0.f, 0.f, 0.f, 1.f };
  memcpy(pose, m, sizeof(m));
}

void orc_synth_frame(uint64_t seed, int frame, int w, int h, const float pose[16], orc_point* out)
{
  const double fx = 0.9 * (double)w;
  const double r00 = pose[0], r01 = pose[1], r10 = pose[4], r11 = pose[5];
  const double ox = pose[3], oy = pose[7], s0 = pose[11];
  const uint64_t fkey = seed ^ ((uint64_t)frame * 0x9E3779B97F4A7C15ULL);
  uint32_t nan_bits = 0x7FC00000u;
  float qnan;
  memcpy(&qnan, &nan_bits, 4);
  for (int v = 0; v < h; v++)
    for (int u = 0; u < w; u++)
    {
      const uint64_t pix = (uint64_t)v * (uint64_t)w + (uint64_t)u;
      const uint64_t r = orc_splitmix64(fkey ^ pix);
      const double dx = (((double)u + 0.5) - 0.5 * (double)w) / fx;
      const double dy = (((double)v + 0.5) - 0.5 * (double)h) / fx;
      const double ax = r00 * dx + r01 * dy; /* world-frame ray slope in x, y (z slope is -1) */
      const double ay = r10 * dx + r11 * dy;
      const double xw0 = ox + s0 * ax, yw0 = oy + s0 * ay;
      const double bump = (0.02 * orc_psin(8.0 * xw0)) * orc_pcos(6.0 * yw0);
      const double s = s0 - bump;
      const double xw = ox + s * ax, yw = oy + s * ay;
      orc_point p;
      memset(&p, 0, sizeof(p));
      p.w = 1.0f;
      if (r % 100u < 5u)
        p.x = p.y = p.z = qnan;
      else
      {
        p.x = (float)(dx * s);
        p.y = (float)(dy * s);
        p.z = (float)s;
      }
      const int64_t stripe = (int64_t)floor(xw * 20.0);
      const int painted = ((stripe & 1) == 0) && (fabs(yw) < 0.06);
      if (painted)
      {
        p.b = (uint8_t)(140 + (int)((r >> 32) % 17u) - 8);
        p.g = (uint8_t)(90 + (int)((r >> 40) % 17u) - 8);
        p.r = (uint8_t)(50 + (int)((r >> 48) % 17u) - 8);
      }
      else
      {
        p.b = (uint8_t)(r >> 8);
        p.g = (uint8_t)(r >> 16);
        p.r = (uint8_t)(r >> 24);
      }
      p.a = 255;
      out[pix] = p;
    }
}
