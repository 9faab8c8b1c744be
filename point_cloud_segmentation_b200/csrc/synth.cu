// synth.cu -- synthetic organized RGB-D scan generator (test / benchmark input).
//
// The reference ships no scan data (.MISSING_LARGE_BLOBS:1-3), so benchmark frames are generated directly in HBM.
// The arithmetic mirrors oracle/synth.c operation for operation and THIS FILE IS COMPILED WITH -fmad=false, so a frame
// generated here is bit-identical to the oracle's: tests/test_synth.py checks that on the GPU.
#include "common.cuh"

namespace pcs
{
__device__ __forceinline__ unsigned long long splitmix64(unsigned long long x)
{
  x += 0x9E3779B97F4A7C15ULL;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ULL;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBULL;
  return x ^ (x >> 31);
}

__device__ __forceinline__ double psin(double x)
{
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

__device__ __forceinline__ double pcos(double x) { return psin(x + 1.5707963267948966); }

__global__ void __launch_bounds__(256) synth_kernel(unsigned long long seed, int first_frame, int w, int h,
                                                   const float* __restrict__ poses, Point8* __restrict__ out)
{
  const int frame_local = blockIdx.y;
  const int frame = first_frame + frame_local;
  const float* pose = poses + (size_t)frame_local * 16;
  const unsigned long long ppf = (unsigned long long)w * h;
  const double fx = 0.9 * (double)w;
  const double r00 = pose[0], r01 = pose[1], r10 = pose[4], r11 = pose[5];
  const double ox = pose[3], oy = pose[7], s0 = pose[11];
  const unsigned long long fkey = seed ^ ((unsigned long long)frame * 0x9E3779B97F4A7C15ULL);
  for (unsigned long long pix = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; pix < ppf;
       pix += (unsigned long long)gridDim.x * blockDim.x)
  {
    const int v = (int)(pix / (unsigned long long)w), u = (int)(pix - (unsigned long long)v * w);
    const unsigned long long r = splitmix64(fkey ^ pix);
    const double dx = (((double)u + 0.5) - 0.5 * (double)w) / fx;
    const double dy = (((double)v + 0.5) - 0.5 * (double)h) / fx;
    const double ax = r00 * dx + r01 * dy;
    const double ay = r10 * dx + r11 * dy;
    const double xw0 = ox + s0 * ax, yw0 = oy + s0 * ay;
    const double bump = (0.02 * psin(8.0 * xw0)) * pcos(6.0 * yw0);
    const double s = s0 - bump;
    const double xw = ox + s * ax, yw = oy + s * ay;
    Point8 p;
    p.w = 1.0f;
    p.p0 = p.p1 = p.p2 = 0u;
    if (r % 100ull < 5ull)
      p.x = p.y = p.z = __uint_as_float(0x7FC00000u);
    else
    {
      p.x = (float)(dx * s);
      p.y = (float)(dy * s);
      p.z = (float)s;
    }
    const long long stripe = (long long)floor(xw * 20.0);
    const bool painted = ((stripe & 1) == 0) && (fabs(yw) < 0.06);
    unsigned int b, g, rr;
    if (painted)
    {
      b = (unsigned int)(140 + (int)((r >> 32) % 17ull) - 8);
      g = (unsigned int)(90 + (int)((r >> 40) % 17ull) - 8);
      rr = (unsigned int)(50 + (int)((r >> 48) % 17ull) - 8);
    }
    else
    {
      b = (unsigned int)((r >> 8) & 0xff);
      g = (unsigned int)((r >> 16) & 0xff);
      rr = (unsigned int)((r >> 24) & 0xff);
    }
    p.rgba = (b & 0xff) | ((g & 0xff) << 8) | ((rr & 0xff) << 16) | (255u << 24);
    st_point_stream(out + (unsigned long long)frame_local * ppf + pix, p);
  }
}
}  // namespace pcs

extern "C" int pcs_synth_frames_dev(pcs_ctx* ctx, uint64_t seed, int first_frame, int n_frames, int width, int height,
                                    const float* poses_dev, pcs_point_xyzrgb* out_dev)
{
  PCS_TRY(pcs::ctx_bind(ctx));
  PCS_CHECK_ARG(poses_dev && out_dev, "NULL argument");
  PCS_CHECK_ARG(n_frames >= 0 && n_frames <= 65535 && width > 0 && height > 0, "bad frame shape");
  if (n_frames == 0)
    return PCS_OK;
  const unsigned long long ppf = (unsigned long long)width * height;
  unsigned long long bx = (ppf + 255) / 256;
  if (bx > 4096)
    bx = 4096;
  dim3 grid((unsigned int)bx, (unsigned int)n_frames);
  pcs::synth_kernel<<<grid, 256, 0, ctx->stream>>>(seed, first_frame, width, height, poses_dev, (pcs::Point8*)out_dev);
  pcs::count_launch(ctx);
  return pcs::check_launch("synth_kernel");
}
