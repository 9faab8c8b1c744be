/*
 * oracle/hsv.c -- TEST INFRASTRUCTURE (see pcs_oracle.h).  CPU restatement of the HSV-threshold detector.
 *
 * Follows pcs_detection/include/pcs_detection/hsv_thresholding.h:45-70:
 *   cv::cvtColor(input, hsv, CV_BGR2HSV)  (8-bit, hue range 180)           :55
 *   cv::inRange(hsv, (95,95,40), (115,220,150), mask)  inclusive both ends  :56-59
 *   inverted -> 255 - mask                                                  :61-62
 * cvtColor/inRange live in OpenCV 3 (pcs_ros/CMakeLists.txt:19), not in the reference tree; the integer
 * algorithm restated here is OpenCV's RGB2HSV_b fixed-point path (hsv_shift = 12, rounded reciprocal
 * tables) and is pinned against cv2 on all 2^24 colours by tests/test_oracle_hsv.py.
 *
 * The optional 5x5 opening has no counterpart in hsvThresholdingDetector; its semantics are those of
 * cv2.morphologyEx(MORPH_OPEN, ones(5,5)) as used by
 * pcs_detection/src_python/pcs_detection/models/histogram_backprojection.py:40-41 (border pixels do not
 * constrain the min / max).
 */
#include "pcs_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

void orc_hsv_default_params(orc_hsv_params* p)
{
  /* hsv_thresholding.h:47-52 */
  p->h_lo = 95;
  p->h_hi = 115;
  p->s_lo = 95;
  p->s_hi = 220;
  p->v_lo = 40;
  p->v_hi = 150;
  p->inverted = 0;
  p->morph_open = 0;
}

void orc_hsv_tables(int32_t sdiv[256], int32_t hdiv[256])
{
  /* OpenCV: sdiv_table[i] = saturate_cast<int>((255 << 12) / (1. * i)); hdiv likewise with 180/(6 i).
   * saturate_cast<int>(double) is cvRound = round half to even (lrint under the default rounding mode). */
  sdiv[0] = 0;
  hdiv[0] = 0;
  for (int i = 1; i < 256; i++)
  {
    sdiv[i] = (int32_t)lrint((double)(255 << 12) / (1.0 * i));
    hdiv[i] = (int32_t)lrint((double)(180 << 12) / (6.0 * i));
  }
}

static inline void hsv_pixel(const int32_t* sdiv, const int32_t* hdiv, int b, int g, int r, int* ho, int* so, int* vo)
{
  int v = b > g ? b : g;
  if (r > v)
    v = r;
  int vmin = b < g ? b : g;
  if (r < vmin)
    vmin = r;
  const int diff = v - vmin;
  const int s = (diff * sdiv[v] + (1 << 11)) >> 12;
  int h;
  if (v == r)
    h = g - b;
  else if (v == g)
    h = (b - r) + 2 * diff;
  else
    h = (r - g) + 4 * diff;
  h = (h * hdiv[diff] + (1 << 11)) >> 12; /* arithmetic shift on a possibly negative value, as OpenCV */
  if (h < 0)
    h += 180;
  *ho = h;
  *so = s;
  *vo = v;
}

void orc_bgr2hsv(const uint8_t* bgr, int w, int h, size_t stride, uint8_t* hsv, size_t hsv_stride)
{
  int32_t sdiv[256], hdiv[256];
  orc_hsv_tables(sdiv, hdiv);
  for (int y = 0; y < h; y++)
  {
    const uint8_t* src = bgr + (size_t)y * stride;
    uint8_t* dst = hsv + (size_t)y * hsv_stride;
    for (int x = 0; x < w; x++)
    {
      int hh, ss, vv;
      hsv_pixel(sdiv, hdiv, src[3 * x], src[3 * x + 1], src[3 * x + 2], &hh, &ss, &vv);
      dst[3 * x] = (uint8_t)hh;
      dst[3 * x + 1] = (uint8_t)ss;
      dst[3 * x + 2] = (uint8_t)vv;
    }
  }
}

static void erode_or_dilate5(const uint8_t* in, int w, int h, uint8_t* out, int dilate)
{
  for (int y = 0; y < h; y++)
    for (int x = 0; x < w; x++)
    {
      int acc = dilate ? 0 : 255;
      for (int dy = -2; dy <= 2; dy++)
      {
        const int yy = y + dy;
        if (yy < 0 || yy >= h)
          continue; /* BORDER_CONSTANT with morphologyDefaultBorderValue: ignored */
        for (int dx = -2; dx <= 2; dx++)
        {
          const int xx = x + dx;
          if (xx < 0 || xx >= w)
            continue;
          const int v = in[(size_t)yy * w + xx];
          if (dilate ? (v > acc) : (v < acc))
            acc = v;
        }
      }
      out[(size_t)y * w + x] = (uint8_t)acc;
    }
}

void orc_morph_open5(const uint8_t* in, int w, int h, uint8_t* out)
{
  uint8_t* tmp = (uint8_t*)malloc((size_t)w * h);
  erode_or_dilate5(in, w, h, tmp, 0);
  erode_or_dilate5(tmp, w, h, out, 1);
  free(tmp);
}

void orc_hsv_mask(const uint8_t* bgr, int w, int h, size_t stride, const orc_hsv_params* p, uint8_t* mask,
                  size_t mask_stride)
{
  /* hsv_thresholding.h:54-59: the reference materialises the HSV image, then thresholds it. */
  uint8_t* hsv = (uint8_t*)malloc((size_t)w * h * 3);
  orc_bgr2hsv(bgr, w, h, stride, hsv, (size_t)w * 3);
  uint8_t* m = (uint8_t*)malloc((size_t)w * h);
  for (size_t i = 0; i < (size_t)w * h; i++)
  {
    const int hh = hsv[3 * i], ss = hsv[3 * i + 1], vv = hsv[3 * i + 2];
    const int in = hh >= p->h_lo && hh <= p->h_hi && ss >= p->s_lo && ss <= p->s_hi && vv >= p->v_lo && vv <= p->v_hi;
    m[i] = in ? 255 : 0;
  }
  if (p->morph_open)
  {
    uint8_t* o = (uint8_t*)malloc((size_t)w * h);
    orc_morph_open5(m, w, h, o);
    free(m);
    m = o;
  }
  for (int y = 0; y < h; y++)
    for (int x = 0; x < w; x++)
    {
      const uint8_t v = m[(size_t)y * w + x];
      mask[(size_t)y * mask_stride + x] = p->inverted ? (uint8_t)(255 - v) : v; /* :61-62 */
    }
  free(m);
  free(hsv);
}

uint64_t orc_hsv_exhaustive_hash(const orc_hsv_params* p, uint64_t* n_positive)
{
  int32_t sdiv[256], hdiv[256];
  orc_hsv_tables(sdiv, hdiv);
  uint64_t hash = 0xcbf29ce484222325ULL, npos = 0;
  for (int r = 0; r < 256; r++)
    for (int g = 0; g < 256; g++)
      for (int b = 0; b < 256; b++)
      {
        int hh, ss, vv;
        hsv_pixel(sdiv, hdiv, b, g, r, &hh, &ss, &vv);
        int in = hh >= p->h_lo && hh <= p->h_hi && ss >= p->s_lo && ss <= p->s_hi && vv >= p->v_lo && vv <= p->v_hi;
        uint8_t m = in ? 255 : 0;
        if (p->inverted)
          m = (uint8_t)(255 - m);
        npos += (m != 0);
        hash = (hash ^ m) * 0x100000001b3ULL;
      }
  if (n_positive)
    *n_positive = npos;
  return hash;
}
