// hsv.cuh -- device-side HSV threshold shared by the image detector kernel and the fused annotate kernel.
//
// Integer algorithm of OpenCV's 8-bit BGR2HSV (hue range 180) followed by cv::inRange, i.e. what
// hsvThresholdingDetector computes (pcs_detection/include/pcs_detection/hsv_thresholding.h:54-59):
//   v = max(b,g,r); diff = v - min(b,g,r); s = (diff*sdiv[v] + 2048) >> 12
//   h = (v==r ? g-b : v==g ? b-r+2diff : r-g+4diff); h = (h*hdiv[diff] + 2048) >> 12; if (h<0) h += 180
// sdiv[i] = round_half_even((255<<12)/i), hdiv[i] = round_half_even((180<<12)/(6i)) -- evaluated here in exact
// integer arithmetic at compile time (OpenCV rounds a double quotient with cvRound; the two agree because no
// quotient lies within 1e-3 of a tie unless it is an exact tie).
#pragma once

#include "common.cuh"

namespace pcs
{
struct HsvLut
{
  int sdiv[256];
  int hdiv[256];
};

constexpr int round_half_even_div(long long n, long long d)
{
  const long long q = n / d, r = n % d;
  if (2 * r > d)
    return (int)(q + 1);
  if (2 * r == d)
    return (int)(q + (q & 1));
  return (int)q;
}

constexpr HsvLut make_hsv_lut()
{
  HsvLut t{};
  t.sdiv[0] = 0;
  t.hdiv[0] = 0;
  for (int i = 1; i < 256; i++)
  {
    t.sdiv[i] = round_half_even_div(255LL << 12, i);
    t.hdiv[i] = round_half_even_div(180LL << 12, 6LL * i);
  }
  return t;
}

static __device__ const HsvLut g_hsv_lut = make_hsv_lut();

// copy both tables (2 KB) into shared memory; call from all threads of the block, then __syncthreads()
__device__ __forceinline__ void load_hsv_lut(int* s_lut /* [512] */)
{
  const int* src = reinterpret_cast<const int*>(&g_hsv_lut);
  for (int i = threadIdx.x; i < 512; i += blockDim.x)
    s_lut[i] = src[i];
}

struct HsvBounds
{
  int h_lo, h_hi, s_lo, s_hi, v_lo, v_hi;
};

// hue [0,179] and saturation [0,255] of a pixel (the two channels the histogram detector back-projects)
__device__ __forceinline__ void bgr_to_hs(const int* __restrict__ s_lut, int b, int g, int r, int& h, int& s)
{
  const int v = max(max(b, g), r);
  const int vmin = min(min(b, g), r);
  const int diff = v - vmin;
  s = (diff * s_lut[v] + 2048) >> 12;
  h = (v == r) ? (g - b) : ((v == g) ? (b - r + 2 * diff) : (r - g + 4 * diff));
  h = (h * s_lut[256 + diff] + 2048) >> 12;
  h += (h < 0) ? 180 : 0;
}

// returns true when (h,s,v) of the pixel lies inside the inclusive bounds
__device__ __forceinline__ bool hsv_in_range(const int* __restrict__ s_lut, int b, int g, int r, const HsvBounds& p)
{
  const int v = max(max(b, g), r);
  const int vmin = min(min(b, g), r);
  const int diff = v - vmin;
  const int s = (diff * s_lut[v] + 2048) >> 12;
  int h = (v == r) ? (g - b) : ((v == g) ? (b - r + 2 * diff) : (r - g + 4 * diff));
  h = (h * s_lut[256 + diff] + 2048) >> 12;
  h += (h < 0) ? 180 : 0;
  return (h >= p.h_lo) & (h <= p.h_hi) & (s >= p.s_lo) & (s <= p.s_hi) & (v >= p.v_lo) & (v <= p.v_hi);
}

// Decision tables of the threshold for one set of bounds, built on the host by exact enumeration (hsv.cu,
// build_decision_tables): a pixel then costs max3 / min3, the hue numerator, two table reads and four compares.
struct HsvDecisionTables
{
  uint32_t vs[256];  // first diff | count << 16
  int2 hx[256];      // each component: (uint16) first X (two's complement) | count << 16
};


__device__ __forceinline__ uint32_t decide(const uint32_t* __restrict__ s_vs, const int2* __restrict__ s_hx, int b, int g,
                                           int r)
{
  const int v = max(max(b, g), r);
  const int diff = v - min(min(b, g), r);
  const int X = (v == r) ? (g - b) : ((v == g) ? (b - r + 2 * diff) : (r - g + 4 * diff));
  const uint32_t e = s_vs[v];
  const int2 hh = s_hx[diff];
  const bool ok_vs = (uint32_t)(diff - (int)(e & 0xffffu)) < (e >> 16);
  const bool ok_h = ((uint32_t)(X - (int)(short)(hh.x & 0xffff)) < ((uint32_t)hh.x >> 16)) |
                    ((uint32_t)(X - (int)(short)(hh.y & 0xffff)) < ((uint32_t)hh.y >> 16));
  return (ok_vs & ok_h) ? 0xffu : 0u;
}


// device copy of the decision tables for `p`, cached in the context; *out = NULL if `p` cannot be tabulated
int hsv_decision_tables(pcs_ctx* ctx, const pcs_hsv_params& p, const HsvDecisionTables** out);
}  // namespace pcs
