/*
 * oracle/backproject.c -- TEST INFRASTRUCTURE (see pcs_oracle.h).  CPU restatement of the histogram back-projection
 * detector, the reference's second /perform_detection backend
 * (pcs_detection/src_python/pcs_detection/models/histogram_backprojection.py:22-52, served by
 * pcs_ros/src/histogram_backprojection_node:29-44).  Every step is an OpenCV call in the reference; the arithmetic
 * restated here is pinned against cv2 by tests/golden/backproject_cv2.npz:
 *   cvtColor(BGR2HSV)                                  :29   (oracle/hsv.c)
 *   calcBackProject([hsv],[0,1],hist,[0,180,0,255],1)  :30   bin = floor(v * bins / (hi - lo)) through a 256-entry table in
 *                                                            double, out-of-range bin -> 0, value = round-half-even(hist)
 *                                                            saturated to 8 bits
 *   filter2D(dst,-1,ellipse 5x5)                       :33-34 17-tap sum, BORDER_REFLECT_101, saturated to 8 bits
 *   threshold(dst, 30, 150, THRESH_BINARY)             :37   v > 30 ? 150 : 0
 *   morphologyEx(MORPH_OPEN, ones(5,5))                :40-41 borders do not constrain min / max
 *   normalize(0,1,NORM_MINMAX) * 255, merge x3         :44-50 a constant image (max == min) normalises to all zeros
 */
#include "pcs_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

void orc_backproject_tables(int h_bins, int s_bins, int32_t tab_h[256], int32_t tab_s[256])
{
  /* OpenCV calcHistLookupTables_8u, uniform ranges: a = bins / (hi - lo), b = -a * lo, idx = floor(j * a + b) */
  const double ah = (double)h_bins / (180.0 - 0.0), bh = -ah * 0.0;
  const double as = (double)s_bins / (255.0 - 0.0), bs = -as * 0.0;
  for (int j = 0; j < 256; j++)
  {
    const int ih = (int)floor(j * ah + bh), is = (int)floor(j * as + bs);
    tab_h[j] = ((unsigned)ih < (unsigned)h_bins) ? ih : -1;
    tab_s[j] = ((unsigned)is < (unsigned)s_bins) ? is : -1;
  }
}

static int reflect101(int p, int len)
{
  if (len == 1)
    return 0;
  while (p < 0 || p >= len)
    p = p < 0 ? -p : 2 * len - 2 - p;
  return p;
}

void orc_histogram_backprojection(const uint8_t* bgr, int w, int h, size_t stride, const float* hist, int h_bins,
                                  int s_bins, int thr_min, int thr_value, uint8_t* out_bgr3)
{
  const size_t n = (size_t)w * h;
  uint8_t* hsv = (uint8_t*)malloc(n * 3);
  uint8_t* bp = (uint8_t*)malloc(n);
  uint8_t* th = (uint8_t*)malloc(n);
  uint8_t* op = (uint8_t*)malloc(n);
  int32_t tab_h[256], tab_s[256];
  orc_backproject_tables(h_bins, s_bins, tab_h, tab_s);
  orc_bgr2hsv(bgr, w, h, stride, hsv, (size_t)w * 3);
  for (size_t i = 0; i < n; i++)
  {
    const int ih = tab_h[hsv[3 * i]], is = tab_s[hsv[3 * i + 1]];
    int v = 0;
    if (ih >= 0 && is >= 0)
    {
      const long r = lrint((double)hist[(size_t)ih * s_bins + is] * 1.0); /* saturate_cast<uchar>(float * scale) */
      v = r < 0 ? 0 : (r > 255 ? 255 : (int)r);
    }
    bp[i] = (uint8_t)v;
  }
  static const uint8_t disc[5][5] = { { 0, 0, 1, 0, 0 }, { 1, 1, 1, 1, 1 }, { 1, 1, 1, 1, 1 }, { 1, 1, 1, 1, 1 }, { 0, 0, 1, 0, 0 } };
  for (int y = 0; y < h; y++)
    for (int x = 0; x < w; x++)
    {
      int sum = 0;
      for (int dy = -2; dy <= 2; dy++)
        for (int dx = -2; dx <= 2; dx++)
          if (disc[dy + 2][dx + 2])
            sum += bp[(size_t)reflect101(y + dy, h) * w + reflect101(x + dx, w)];
      if (sum > 255)
        sum = 255;
      th[(size_t)y * w + x] = sum > thr_min ? (uint8_t)thr_value : 0;
    }
  orc_morph_open5(th, w, h, op);
  int mn = 255, mx = 0;
  for (size_t i = 0; i < n; i++)
  {
    if (op[i] < mn)
      mn = op[i];
    if (op[i] > mx)
      mx = op[i];
  }
  for (size_t i = 0; i < n; i++)
  {
    /* normalize to [0,1] then * 255: with max == min OpenCV's scale is 0 and everything becomes 0 */
    int v = 0;
    if (mx > mn)
      v = (int)lrint((double)(op[i] - mn) * (1.0 / (double)(mx - mn))) * 255;
    out_bgr3[3 * i] = out_bgr3[3 * i + 1] = out_bgr3[3 * i + 2] = (uint8_t)v;
  }
  free(op);
  free(th);
  free(bp);
  free(hsv);
}
