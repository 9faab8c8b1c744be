// hsv.cu -- stage 1: fused BGR->HSV + inRange (+ optional 5x5 opening) over an 8UC3 image.
//
// Replaces hsvThresholdingDetector (pcs_detection/include/pcs_detection/hsv_thresholding.h:45-70) and the detector
// service body (pcs_ros/src/hsv_thresholding_node.cpp:8-28).  The HSV image the reference materialises never
// exists: each block stages a tile of BGR rows in shared memory with 128-bit loads, thresholds it there, optionally
// erodes/dilates it there, and writes the mask with 128-bit stores.
//
// HBM-bound: 3 B read + 1 B written per pixel (4 B/px algorithmic; +9 % halo re-reads when the opening is on).
#include "common.cuh"
#include "hsv.cuh"

namespace pcs
{
constexpr int kTileW = 128;
constexpr int kTileH = 8;
constexpr int kHsvThreads = 256;

template <int HALO>
struct TileGeom
{
  static constexpr int RW = kTileW + 2 * HALO;  // region width in pixels
  static constexpr int RH = kTileH + 2 * HALO;
  static constexpr int ROW_BYTES = ((RW * 3 + 15) / 16) * 16 + 32;  // room for the 16-byte phase of the source
};

template <bool MORPH, int OUT_CH>
__global__ void __launch_bounds__(kHsvThreads) hsv_mask_kernel(const uint8_t* __restrict__ bgr, int width, int height,
                                                              size_t stride, HsvBounds bounds, int inverted,
                                                              uint8_t* __restrict__ mask, size_t mask_stride)
{
  constexpr int HALO = MORPH ? 4 : 0;
  using G = TileGeom<HALO>;
  __shared__ int s_lut[512];
  __shared__ __align__(16) uint8_t s_bgr[G::RH][G::ROW_BYTES];
  __shared__ uint8_t s_m[G::RH][G::RW + 4];
  __shared__ uint8_t s_e[MORPH ? (kTileH + 4) : 1][MORPH ? (kTileW + 8) : 4];

  load_hsv_lut(s_lut);

  const int x0 = blockIdx.x * kTileW - HALO;  // region origin (may be negative)
  const int y0 = blockIdx.y * kTileH - HALO;
  const int xs = max(x0, 0), xe = min(x0 + G::RW, width);  // valid pixel columns [xs, xe)

  // ---- stage BGR rows: 16-byte chunks where the global address allows, bytes at the ragged ends ----------------
  for (int r = threadIdx.x / 32; r < G::RH; r += kHsvThreads / 32)
  {
    const int y = y0 + r;
    if (y < 0 || y >= height)
      continue;
    const uint8_t* row = bgr + (size_t)y * stride;
    const uintptr_t beg = (uintptr_t)(row + 3 * (size_t)xs), end = (uintptr_t)(row + 3 * (size_t)xe);
    const int phase = (int)(beg & 15u);
    uint8_t* dst = &s_bgr[r][0];  // dst[phase + k] holds byte beg + k
    const uintptr_t a0 = (beg + 15u) & ~(uintptr_t)15u, a1 = end & ~(uintptr_t)15u;
    const int lane = threadIdx.x & 31;
    if (a0 <= a1)
    {
      for (uintptr_t a = beg + lane; a < a0; a += 32)
        dst[phase + (int)(a - beg)] = *(const uint8_t*)a;
      for (uintptr_t a = a0 + 16u * lane; a < a1; a += 16u * 32u)
        *reinterpret_cast<uint4*>(dst + phase + (a - beg)) = __ldg(reinterpret_cast<const uint4*>(a));
      for (uintptr_t a = a1 + lane; a < end; a += 32)
        dst[phase + (int)(a - beg)] = *(const uint8_t*)a;
    }
    else
      for (uintptr_t a = beg + lane; a < end; a += 32)
        dst[phase + (int)(a - beg)] = *(const uint8_t*)a;
  }
  __syncthreads();

  // ---- threshold the region -------------------------------------------------------------------------------------
  for (int i = threadIdx.x; i < G::RH * G::RW; i += kHsvThreads)
  {
    const int r = i / G::RW, c = i - r * G::RW;
    const int x = x0 + c, y = y0 + r;
    uint8_t m = 255;  // outside the image: does not constrain the erosion
    if (x >= 0 && x < width && y >= 0 && y < height)
    {
      const uint8_t* row = bgr + (size_t)y * stride;
      const int phase = (int)((uintptr_t)(row + 3 * (size_t)xs) & 15u);
      const uint8_t* px = &s_bgr[r][phase + 3 * (x - xs)];
      m = hsv_in_range(s_lut, px[0], px[1], px[2], bounds) ? 255 : 0;
    }
    s_m[r][c] = m;
  }
  __syncthreads();

  if (MORPH)
  {
    // erosion over the tile grown by 2; positions outside the image become 0 so the dilation ignores them
    for (int i = threadIdx.x; i < (kTileH + 4) * (kTileW + 4); i += kHsvThreads)
    {
      const int r = i / (kTileW + 4), c = i - r * (kTileW + 4);
      const int x = x0 + 2 + c, y = y0 + 2 + r;
      uint8_t e = 0;
      if (x >= 0 && x < width && y >= 0 && y < height)
      {
        e = 255;
#pragma unroll
        for (int dy = 0; dy < 5; dy++)
#pragma unroll
          for (int dx = 0; dx < 5; dx++)
            e = min(e, s_m[r + dy][c + dx]);
      }
      s_e[r][c] = e;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < kTileH * kTileW; i += kHsvThreads)
    {
      const int r = i / kTileW, c = i - r * kTileW;
      uint8_t d = 0;
#pragma unroll
      for (int dy = 0; dy < 5; dy++)
#pragma unroll
        for (int dx = 0; dx < 5; dx++)
          d = max(d, s_e[r + dy][c + dx]);
      s_bgr[r][c] = d;  // reuse the BGR staging area as the output tile (row pitch ROW_BYTES >= kTileW)
    }
    __syncthreads();
  }

  // ---- write the tile ---------------------------------------------------------------------------------------------
  const int tx0 = blockIdx.x * kTileW, ty0 = blockIdx.y * kTileH;
  const uint8_t inv = inverted ? 255 : 0;
  if (OUT_CH == 1)
  {
    // one 16-pixel group per thread-iteration
    for (int i = threadIdx.x; i < kTileH * (kTileW / 16); i += kHsvThreads)
    {
      const int r = i / (kTileW / 16), cg = (i - r * (kTileW / 16)) * 16;
      const int y = ty0 + r, x = tx0 + cg;
      if (y >= height || x >= width)
        continue;
      uint8_t v[16];
#pragma unroll
      for (int k = 0; k < 16; k++)
        v[k] = (MORPH ? s_bgr[r][cg + k] : s_m[r][cg + k]) ^ inv;
      uint8_t* out = mask + (size_t)y * mask_stride + x;
      if (x + 16 <= width && (((uintptr_t)out) & 15u) == 0)
      {
        uint4 q;
        memcpy(&q, v, 16);
        *reinterpret_cast<uint4*>(out) = q;
      }
      else
        for (int k = 0; k < 16 && x + k < width; k++)
          out[k] = v[k];
    }
  }
  else
  {
    for (int i = threadIdx.x; i < kTileH * kTileW; i += kHsvThreads)
    {
      const int r = i / kTileW, c = i - r * kTileW;
      const int y = ty0 + r, x = tx0 + c;
      if (y >= height || x >= width)
        continue;
      const uint8_t v = (MORPH ? s_bgr[r][c] : s_m[r][c]) ^ inv;
      uint8_t* out = mask + (size_t)y * mask_stride + 3 * (size_t)x;
      out[0] = v;
      out[1] = v;
      out[2] = v;
    }
  }
}

int launch_hsv_mask(pcs_ctx* ctx, const uint8_t* bgr_dev, int width, int height, size_t stride,
                    const pcs_hsv_params* p, uint8_t* mask_dev, size_t mask_stride, int out_channels)
{
  HsvBounds b = { p->h_lo, p->h_hi, p->s_lo, p->s_hi, p->v_lo, p->v_hi };
  dim3 grid((unsigned)((width + kTileW - 1) / kTileW), (unsigned)((height + kTileH - 1) / kTileH));
  dim3 block(kHsvThreads);
  const int inv = p->inverted ? 1 : 0;
  if (p->morph_open)
  {
    if (out_channels == 1)
      hsv_mask_kernel<true, 1><<<grid, block, 0, ctx->stream>>>(bgr_dev, width, height, stride, b, inv, mask_dev, mask_stride);
    else
      hsv_mask_kernel<true, 3><<<grid, block, 0, ctx->stream>>>(bgr_dev, width, height, stride, b, inv, mask_dev, mask_stride);
  }
  else
  {
    if (out_channels == 1)
      hsv_mask_kernel<false, 1><<<grid, block, 0, ctx->stream>>>(bgr_dev, width, height, stride, b, inv, mask_dev, mask_stride);
    else
      hsv_mask_kernel<false, 3><<<grid, block, 0, ctx->stream>>>(bgr_dev, width, height, stride, b, inv, mask_dev, mask_stride);
  }
  count_launch(ctx);
  return check_launch("hsv_mask_kernel");
}
}  // namespace pcs

extern "C" {

int pcs_hsv_mask_dev(pcs_ctx* ctx, const uint8_t* bgr_dev, int width, int height, size_t stride,
                     const pcs_hsv_params* params, uint8_t* mask_dev, size_t mask_stride)
{
  PCS_TRY(pcs::ctx_bind(ctx));
  PCS_CHECK_ARG(bgr_dev && mask_dev && params, "NULL argument");
  PCS_CHECK_ARG(width > 0 && height > 0, "empty image");
  PCS_CHECK_ARG(stride >= (size_t)width * 3 && mask_stride >= (size_t)width, "row pitch smaller than a row");
  return pcs::launch_hsv_mask(ctx, bgr_dev, width, height, stride, params, mask_dev, mask_stride, 1);
}

static int hsv_host(pcs_ctx* ctx, const uint8_t* bgr, int width, int height, size_t stride, const pcs_hsv_params* params,
                    uint8_t* out, size_t out_stride, int channels)
{
  PCS_TRY(pcs::ctx_bind(ctx));
  PCS_CHECK_ARG(bgr && out && params, "NULL argument");
  PCS_CHECK_ARG(width > 0 && height > 0, "empty image");
  PCS_CHECK_ARG(stride >= (size_t)width * 3 && out_stride >= (size_t)width * channels, "row pitch smaller than a row");
  void *d_in = nullptr, *d_out = nullptr;
  // the last row of a pitched image only owns width * channels bytes
  const size_t in_bytes = stride * (size_t)(height - 1) + (size_t)width * 3;
  const size_t out_bytes = out_stride * (size_t)(height - 1) + (size_t)width * channels;
  PCS_TRY(pcs::ctx_staging(ctx, 0, in_bytes, &d_in));
  PCS_TRY(pcs::ctx_staging(ctx, 1, out_bytes, &d_out));
  PCS_CUDA(cudaMemcpyAsync(d_in, bgr, in_bytes, cudaMemcpyHostToDevice, ctx->stream));
  if (out_stride != (size_t)width * channels)
    PCS_CUDA(cudaMemcpyAsync(d_out, out, out_bytes, cudaMemcpyHostToDevice, ctx->stream));  // keep caller's padding
  PCS_TRY(pcs::launch_hsv_mask(ctx, (const uint8_t*)d_in, width, height, stride, params, (uint8_t*)d_out, out_stride,
                               channels));
  PCS_CUDA(cudaMemcpyAsync(out, d_out, out_bytes, cudaMemcpyDeviceToHost, ctx->stream));
  PCS_CUDA(cudaStreamSynchronize(ctx->stream));
  ctx->stats.bytes_h2d += in_bytes;
  ctx->stats.bytes_d2h += out_bytes;
  return PCS_OK;
}

int pcs_hsv_mask(pcs_ctx* ctx, const uint8_t* bgr, int width, int height, size_t stride, const pcs_hsv_params* params,
                 uint8_t* mask, size_t mask_stride)
{
  return hsv_host(ctx, bgr, width, height, stride, params, mask, mask_stride, 1);
}

int pcs_hsv_annotation_image(pcs_ctx* ctx, const uint8_t* bgr, int width, int height, size_t stride,
                             const pcs_hsv_params* params, uint8_t* annotation_bgr)
{
  return hsv_host(ctx, bgr, width, height, stride, params, annotation_bgr, (size_t)width * 3, 3);
}

}  // extern "C"
