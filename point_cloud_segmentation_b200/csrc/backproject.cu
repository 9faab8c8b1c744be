// backproject.cu -- histogram back-projection detector: the reference's second /perform_detection backend,
// HistogramBackprojection.annotate_image (pcs_detection/src_python/pcs_detection/models/histogram_backprojection.py:22-52,
// served by pcs_ros/src/histogram_backprojection_node:29-44).
//
//   cvtColor(BGR2HSV) + calcBackProject on (H,S)   -> backproject_kernel   (hue/saturation bin tables + histogram lookup)
//   filter2D(ellipse 5x5) + threshold(30 -> 150)   -> disc_threshold_kernel (17-tap sum, BORDER_REFLECT_101)
//   morphologyEx(MORPH_OPEN, ones 5x5)             -> erode5_kernel, dilate5_kernel (borders do not constrain)
//   normalize(0,1,MINMAX) * 255, merge x3          -> finalize_kernel (needs the global min / max: two flags)
// Integer arithmetic throughout; bit-exact against cv2 (tests/golden/backproject_cv2.npz).  Plain global-memory passes:
// this detector is an alternative front end, not the fused hot path.
#include <math.h>

#include "common.cuh"
#include "hsv.cuh"

namespace pcs
{
struct BinTables
{
  short h[256];
  short s[256];
};

__global__ void backproject_kernel(const uint8_t* __restrict__ bgr, int w, int h, size_t stride, BinTables tabs,
                                   const float* __restrict__ hist, int s_bins, uint8_t* __restrict__ out)
{
  __shared__ int s_lut[512];
  load_hsv_lut(s_lut);
  __syncthreads();
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
  if (x >= w)
    return;
  const uint8_t* px = bgr + (size_t)y * stride + 3 * (size_t)x;
  int hh, ss;
  bgr_to_hs(s_lut, px[0], px[1], px[2], hh, ss);
  const int ih = tabs.h[hh], is = tabs.s[ss];
  int v = 0;
  if (ih >= 0 && is >= 0)
    v = min(max(__float2int_rn(__ldg(hist + (size_t)ih * s_bins + is)), 0), 255);  // saturate_cast<uchar>, half to even
  out[(size_t)y * w + x] = (uint8_t)v;
}

__device__ __forceinline__ int reflect101(int p, int len)
{
  if (len == 1)
    return 0;
  while (p < 0 || p >= len)
    p = p < 0 ? -p : 2 * len - 2 - p;
  return p;
}

__global__ void disc_threshold_kernel(const uint8_t* __restrict__ in, int w, int h, int thr_min, int thr_value,
                                      uint8_t* __restrict__ out)
{
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
  if (x >= w)
    return;
  int sum = 0;
#pragma unroll
  for (int dy = -2; dy <= 2; dy++)
  {
    const int yy = reflect101(y + dy, h);
    const int r = (dy == -2 || dy == 2) ? 0 : 2;  // getStructuringElement(MORPH_ELLIPSE, (5,5)): rows 0 and 4 hold one tap
#pragma unroll
    for (int dx = -2; dx <= 2; dx++)
      if (dx >= -r && dx <= r)
        sum += in[(size_t)yy * w + reflect101(x + dx, w)];
  }
  out[(size_t)y * w + x] = min(sum, 255) > thr_min ? (uint8_t)thr_value : 0;
}

template <bool DILATE>
__global__ void morph5_kernel(const uint8_t* __restrict__ in, int w, int h, uint8_t* __restrict__ out,
                              unsigned int* __restrict__ flags)
{
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
  if (x >= w)
    return;
  int acc = DILATE ? 0 : 255;
#pragma unroll
  for (int dy = -2; dy <= 2; dy++)
  {
    const int yy = y + dy;
    if (yy < 0 || yy >= h)
      continue;
#pragma unroll
    for (int dx = -2; dx <= 2; dx++)
    {
      const int xx = x + dx;
      if (xx < 0 || xx >= w)
        continue;
      const int v = in[(size_t)yy * w + xx];
      acc = DILATE ? max(acc, v) : min(acc, v);
    }
  }
  out[(size_t)y * w + x] = (uint8_t)acc;
  if (DILATE)
  {
    // the min-max normalisation only needs to know whether both values occur
    const unsigned any_set = __any_sync(__activemask(), acc != 0), any_clear = __any_sync(__activemask(), acc == 0);
    if ((threadIdx.x & 31) == 0 || x == 0)
    {
      if (any_set)
        atomicOr(flags, 1u);
      if (any_clear)
        atomicOr(flags, 2u);
    }
  }
}

__global__ void finalize_kernel(const uint8_t* __restrict__ in, size_t n, const unsigned int* __restrict__ flags,
                                uint8_t* __restrict__ out3)
{
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n)
    return;
  // cv::normalize(0, 1, NORM_MINMAX): with max == min the scale is 0 and the whole image becomes 0
  const uint8_t v = (*flags == 3u && in[i] != 0) ? 255 : 0;
  out3[3 * i] = v;
  out3[3 * i + 1] = v;
  out3[3 * i + 2] = v;
}
}  // namespace pcs

using namespace pcs;

extern "C" int pcs_histogram_backprojection(pcs_ctx* ctx, const uint8_t* bgr, int width, int height, size_t stride,
                                            const float* histogram, int h_bins, int s_bins, int threshold_min,
                                            int threshold_value, uint8_t* annotation_bgr)
{
  PCS_TRY(ctx_bind(ctx));
  PCS_CHECK_ARG(bgr && histogram && annotation_bgr, "NULL argument");
  PCS_CHECK_ARG(width > 0 && height > 0 && height <= 65535, "bad image shape");
  PCS_CHECK_ARG(stride >= (size_t)width * 3, "row pitch smaller than a row");
  PCS_CHECK_ARG(h_bins > 0 && s_bins > 0 && h_bins <= 32767 && s_bins <= 32767, "bad histogram shape");
  PCS_CHECK_ARG(threshold_value >= 0 && threshold_value <= 255, "threshold value must fit 8 bits");
  // OpenCV calcHistLookupTables_8u, uniform ranges [0,180) and [0,255): idx = floor(j * bins / (hi - lo)) in double
  BinTables tabs;
  const double ah = (double)h_bins / 180.0, as = (double)s_bins / 255.0;
  for (int j = 0; j < 256; j++)
  {
    const int ih = (int)floor(j * ah + (-ah * 0.0)), is = (int)floor(j * as + (-as * 0.0));
    tabs.h[j] = (short)(((unsigned)ih < (unsigned)h_bins) ? ih : -1);
    tabs.s[j] = (short)(((unsigned)is < (unsigned)s_bins) ? is : -1);
  }
  const size_t n = (size_t)width * height;
  const size_t in_bytes = stride * (size_t)(height - 1) + (size_t)width * 3, hist_bytes = (size_t)h_bins * s_bins * sizeof(float);
  void *d_in = nullptr, *d_tmp = nullptr, *d_out = nullptr;
  auto al = [](size_t b) { return (b + 255) & ~(size_t)255; };
  PCS_TRY(ctx_staging(ctx, 0, al(in_bytes) + al(hist_bytes), &d_in));
  PCS_TRY(ctx_staging(ctx, 1, 2 * al(n) + 256, &d_tmp));
  PCS_TRY(ctx_staging(ctx, 2, n * 3, &d_out));
  float* d_hist = (float*)((char*)d_in + al(in_bytes));
  uint8_t *t0 = (uint8_t*)d_tmp, *t1 = t0 + al(n);
  unsigned int* d_flags = (unsigned int*)(t1 + al(n));
  PCS_CUDA(cudaMemcpyAsync(d_in, bgr, in_bytes, cudaMemcpyHostToDevice, ctx->stream));
  PCS_CUDA(cudaMemcpyAsync(d_hist, histogram, hist_bytes, cudaMemcpyHostToDevice, ctx->stream));
  PCS_CUDA(cudaMemsetAsync(d_flags, 0, sizeof(unsigned int), ctx->stream));
  const dim3 block(128), grid((unsigned)((width + 127) / 128), (unsigned)height);
  backproject_kernel<<<grid, block, 0, ctx->stream>>>((const uint8_t*)d_in, width, height, stride, tabs, d_hist, s_bins, t0);
  disc_threshold_kernel<<<grid, block, 0, ctx->stream>>>(t0, width, height, threshold_min, threshold_value, t1);
  morph5_kernel<false><<<grid, block, 0, ctx->stream>>>(t1, width, height, t0, d_flags);
  morph5_kernel<true><<<grid, block, 0, ctx->stream>>>(t0, width, height, t1, d_flags);
  finalize_kernel<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(t1, n, d_flags, (uint8_t*)d_out);
  count_launch(ctx, 5);
  PCS_TRY(check_launch("histogram back-projection kernels"));
  PCS_CUDA(cudaMemcpyAsync(annotation_bgr, d_out, n * 3, cudaMemcpyDeviceToHost, ctx->stream));
  PCS_CUDA(cudaStreamSynchronize(ctx->stream));
  ctx->stats.bytes_h2d += in_bytes + hist_bytes;
  ctx->stats.bytes_d2h += n * 3;
  return PCS_OK;
}
