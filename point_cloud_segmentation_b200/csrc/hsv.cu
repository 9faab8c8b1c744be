// hsv.cu -- stage 1: fused BGR->HSV + inRange (+ optional 5x5 opening) over an 8UC3 image.
//
// Replaces hsvThresholdingDetector (pcs_detection/include/pcs_detection/hsv_thresholding.h:45-70) and the detector
// service body (pcs_ros/src/hsv_thresholding_node.cpp:8-28).  The HSV image the reference materialises never
// exists: the threshold runs on BGR rows staged with 128-bit loads (dense images: hsv_mask_flat_kernel), the optional
// opening runs on bit-packed masks in shared memory (hsv_open_kernel), and the mask is written with wide stores.
//
// HBM-bound: 3 B read + 1 B written per pixel (4 B/px algorithmic; the opening re-reads a 4-pixel halo through L1/L2).
#include "common.cuh"
#include "hsv.cuh"

namespace pcs
{
constexpr int kTileW = 128;
constexpr int kTileH = 8;
constexpr int kHsvThreads = 256;

struct TileGeom
{
  static constexpr int RW = kTileW;
  static constexpr int RH = kTileH;
  static constexpr int ROW_BYTES = ((RW * 3 + 15) / 16) * 16 + 32;  // room for the 16-byte phase of the source
};

// generic (pitched / unaligned / 3-channel output) threshold kernel, no opening
template <int OUT_CH>
__global__ void __launch_bounds__(kHsvThreads) hsv_mask_kernel(const uint8_t* __restrict__ bgr, int width, int height,
                                                              size_t stride, HsvBounds bounds, int inverted,
                                                              uint8_t* __restrict__ mask, size_t mask_stride)
{
  using G = TileGeom;
  __shared__ int s_lut[512];
  __shared__ __align__(16) uint8_t s_bgr[G::RH][G::ROW_BYTES];
  __shared__ uint8_t s_m[G::RH][G::RW + 4];

  load_hsv_lut(s_lut);

  const int x0 = blockIdx.x * kTileW;
  const int y0 = blockIdx.y * kTileH;
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

  // ---- threshold the tile -----------------------------------------------------------------------------------------
  for (int i = threadIdx.x; i < G::RH * G::RW; i += kHsvThreads)
  {
    const int r = i / G::RW, c = i - r * G::RW;
    const int x = x0 + c, y = y0 + r;
    uint8_t m = 0;
    if (x < width && y < height)
    {
      const uint8_t* row = bgr + (size_t)y * stride;
      const int phase = (int)((uintptr_t)(row + 3 * (size_t)xs) & 15u);
      const uint8_t* px = &s_bgr[r][phase + 3 * (x - xs)];
      m = hsv_in_range(s_lut, px[0], px[1], px[2], bounds) ? 255 : 0;
    }
    s_m[r][c] = m;
  }
  __syncthreads();

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
        v[k] = s_m[r][cg + k] ^ inv;
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
      const uint8_t v = s_m[r][c] ^ inv;
      uint8_t* out = mask + (size_t)y * mask_stride + 3 * (size_t)x;
      out[0] = v;
      out[1] = v;
      out[2] = v;
    }
  }
}

// ---- threshold + 5x5 opening on bit-packed masks ------------------------------------------------------------------------
// A CTA takes a tile of 256 columns x tile_h (<= 64) rows.  Phase A thresholds the tile grown by 4 pixels on every side
// (the opening's 9x9 dependency) and packs the result with warp ballots: 32 pixels per word, 10 words per row (word 0 /
// word 9 carry the 4 halo columns).  Phase B opens the packed rows separably: a 5-wide min is 4 shifts + 4 ANDs of a
// 64-bit window, a 5-tall min is 4 ANDs across lanes that each hold one row (shuffles); dilation likewise with ORs.  Each
// warp owns 32 columns and produces 24 finished rows per step from 32 thresholded rows.  cv2's border rule: pixels
// outside the image take no part = thresholded pixels outside count as 1 for the erosion, eroded pixels outside as 0.
constexpr int kOpenImgTW = 256;
constexpr int kOpenImgMaxTH = 64;
constexpr int kOpenImgWords = kOpenImgTW / 32 + 2;

template <int OUT_CH>
__global__ void __launch_bounds__(kOpenImgTW) hsv_open_kernel(const uint8_t* __restrict__ bgr, int width, int height,
                                                            size_t stride, HsvBounds bounds, int inverted,
                                                            uint8_t* __restrict__ mask, size_t mask_stride, int tile_h)
{
  __shared__ int s_lut[512];
  __shared__ unsigned int s_raw[kOpenImgMaxTH + 8][kOpenImgWords];
  load_hsv_lut(s_lut);
  const unsigned int tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
  const int x0 = blockIdx.x * kOpenImgTW, y0 = blockIdx.y * tile_h;
  const int rows_out = min(tile_h, height - y0);
  const int R = rows_out + 8;  // thresholded rows e = 0 .. R-1 <-> image rows y0 - 4 + e
  __syncthreads();

  // ---- A: threshold + pack ------------------------------------------------------------------------------------------------
  {
    const int col = x0 + (int)tid;
    const bool col_ok = col < width;
    const uint8_t* px0 = bgr + 3 * (size_t)(col_ok ? col : 0);
#pragma unroll 4
    for (int e = 0; e < R; e++)
    {
      const int y = y0 - 4 + e;
      bool bit = true;
      if (col_ok && y >= 0 && y < height)
      {
        const uint8_t* px = px0 + (size_t)y * stride;
        bit = hsv_in_range(s_lut, __ldg(px), __ldg(px + 1), __ldg(px + 2), bounds);
      }
      const unsigned int word = __ballot_sync(0xffffffffu, bit);
      if (lane == 0)
        s_raw[e][1 + warp] = word;
    }
    // halo columns: pair p = 8 * row + halo column, 32 pairs (4 rows) per warp pass
    for (int p = (int)tid; p < ((R * 8 + 31) & ~31); p += kOpenImgTW)
    {
      const int e = p >> 3, hc = p & 7;
      const int y = y0 - 4 + e;
      const int hcol = hc < 4 ? x0 - 4 + hc : x0 + kOpenImgTW + (hc - 4);
      bool bit = true;
      if (e < R && y >= 0 && y < height && hcol >= 0 && hcol < width)
      {
        const uint8_t* px = bgr + (size_t)y * stride + 3 * (size_t)hcol;
        bit = hsv_in_range(s_lut, __ldg(px), __ldg(px + 1), __ldg(px + 2), bounds);
      }
      const unsigned int hb = __ballot_sync(0xffffffffu, bit);
      const int er = ((p & ~31) >> 3) + (int)lane;  // lanes 0..3 store the 4 rows of this pass
      if (lane < 4u && er < R)
      {
        const unsigned int byte = (hb >> (8u * lane)) & 0xffu;
        s_raw[er][0] = (byte & 0xfu) << 28;
        s_raw[er][kOpenImgWords - 1] = byte >> 4;
      }
    }
  }
  __syncthreads();

  // ---- B + C: open 32 columns x 24 rows per step, write them ------------------------------------------------------------------
  unsigned long long col_valid = ~0ull;  // columns [x0 + 32 warp - 16, +64) inside the image
  {
    const int lo = x0 + 32 * (int)warp - 16;
    if (lo < 0)
      col_valid &= ~0ull << (-lo);
    const int n = width - lo;
    if (n < 64)
      col_valid &= (n <= 0) ? 0ull : ((1ull << n) - 1ull);
  }
  const unsigned int inv4 = inverted ? 0xffffffffu : 0u;
  for (int eb = 0; eb < rows_out; eb += 24)
  {
    const unsigned int* rr = s_raw[min(eb + (int)lane, R - 1)];
    const unsigned long long win =
        ((unsigned long long)rr[warp] >> 16) | ((unsigned long long)rr[warp + 1] << 16) | ((unsigned long long)rr[warp + 2] << 48);
    // 5-wide min in three shift/AND steps: pairs, triples (bits i .. i+2), then the triples at i-2 and at i
    const unsigned long long w2 = win & (win >> 1), w3 = w2 & (win >> 2);
    const unsigned long long he = w3 & (w3 << 2);
    unsigned long long er5 = he;  // lane l <= 27: eroded row e = eb + 2 + l
#pragma unroll
    for (int d = 1; d <= 4; d++)
      er5 &= __shfl_down_sync(0xffffffffu, he, d);
    const int y_e = y0 - 4 + eb + 2 + (int)lane;
    er5 = (y_e >= 0 && y_e < height) ? (er5 & col_valid) : 0ull;
    const unsigned long long d2 = er5 | (er5 >> 1), d3 = d2 | (er5 >> 2);
    const unsigned long long hd = d3 | (d3 << 2);
    unsigned long long op = hd;  // lane l <= 23: opened row e = eb + 4 + l <-> image row y0 + eb + l
#pragma unroll
    for (int d = 1; d <= 4; d++)
      op |= __shfl_down_sync(0xffffffffu, hd, d);
    const unsigned int fin = (unsigned int)(op >> 16);

    if (OUT_CH == 1)
    {
      // 4 rows per pass: lane = 8 * row + group of 4 pixels -> one 32-bit store (bytes when the address is not aligned)
      const int i = (int)(lane >> 3), g4 = (int)(lane & 7u) * 4;
#pragma unroll
      for (int q = 0; q < 6; q++)
      {
        const int ri = 4 * q + i;
        const unsigned int word = __shfl_sync(0xffffffffu, fin, ri);
        const int y = y0 + eb + ri, x = x0 + 32 * (int)warp + g4;
        if (eb + ri >= rows_out || x >= width)
          continue;
        const unsigned int nib = (word >> g4) & 0xfu;
        const unsigned int bytes = (((nib * 0x00204081u) & 0x01010101u) * 0xffu) ^ inv4;
        uint8_t* out = mask + (size_t)y * mask_stride + x;
        if (x + 4 <= width && (((uintptr_t)out) & 3u) == 0)
          *reinterpret_cast<unsigned int*>(out) = bytes;
        else
          for (int k = 0; k < 4 && x + k < width; k++)
            out[k] = (uint8_t)(bytes >> (8 * k));
      }
    }
    else
    {
#pragma unroll 4
      for (int ri = 0; ri < 24; ri++)
      {
        const unsigned int word = __shfl_sync(0xffffffffu, fin, ri);
        const int y = y0 + eb + ri, x = x0 + (int)tid;
        if (eb + ri >= rows_out || x >= width)
          continue;
        const uint8_t v = (uint8_t)((((word >> lane) & 1u) ? 255u : 0u) ^ (inv4 & 0xffu));
        uint8_t* out = mask + (size_t)y * mask_stride + 3 * (size_t)x;
        out[0] = v;
        out[1] = v;
        out[2] = v;
      }
    }
  }
}

// ---- contiguous fast path -----------------------------------------------------------------------------------------------
// When the image and the mask are dense (row pitch == row size) and 16-byte aligned, the image is a flat array of
// pixels.  A CTA takes 4096 pixels: the 12 KiB of BGR bytes are staged in shared memory with coalesced 128-bit loads
// (lane l fetches bytes [16 l, 16 l + 16) of each 512-byte run), every thread then owns 16 consecutive pixels = three
// 128-bit shared-memory reads at a 48-byte stride (conflict-free: 12 words per lane, gcd(12, 32) = 4 and a quarter-warp
// covers all 32 banks), thresholds them in registers and writes its 16 mask bytes with one 128-bit store, so a warp
// writes 512 contiguous bytes.
//
// The per-pixel arithmetic is folded into two small decision tables built on the host for the current thresholds:
//   vs[v]    = { first diff, number of diffs } for which v and s = (diff * sdiv[v] + 2048) >> 12 are both in range
//              (s is monotone in diff, so the accepted diffs form one run);
//   hx[diff] = up to two runs { first X, count } of the hue numerator X = g-b | b-r+2diff | r-g+4diff for which
//              h = (X * hdiv[diff] + 2048) >> 12 (+180 if negative) is in range (h is monotone in X on either side of
//              the wrap, so an interval of hues accepts at most two runs of X).
// Both are enumerated with the exact reference formulas, so the kernel's decisions are the reference's bit for bit; a
// pixel then costs max3 / min3, the numerator, two table reads and four compares instead of two multiplies, two shifts,
// a wrap and six compares.
constexpr int kFastThreads = 256;
constexpr int kFastPxPerThread = 16;
constexpr int kFastPxPerCta = kFastThreads * kFastPxPerThread;  // 4096

// false if the thresholds cannot be expressed (never the case for an interval of hues; kept as a guard)
static bool build_decision_tables(const pcs_hsv_params& p, HsvDecisionTables& t)
{
  const HsvLut lut = make_hsv_lut();
  for (int v = 0; v < 256; v++)
  {
    int first = -1, last = -2;
    for (int diff = 0; diff <= v; diff++)
    {
      const int sat = (diff * lut.sdiv[v] + 2048) >> 12;
      const bool ok = v >= p.v_lo && v <= p.v_hi && sat >= p.s_lo && sat <= p.s_hi;
      if (ok)
      {
        if (first < 0)
          first = diff;
        else if (diff != last + 1)
          return false;
        last = diff;
      }
    }
    t.vs[v] = first < 0 ? 0u : ((uint32_t)first | ((uint32_t)(last - first + 1) << 16));
  }
  for (int diff = 0; diff < 256; diff++)
  {
    int start[2] = { 0, 0 }, count[2] = { 0, 0 }, n_runs = 0;
    bool in_run = false;
    for (int X = -255; X <= 6 * 255; X++)
    {
      int h = (X * lut.hdiv[diff] + 2048) >> 12;  // arithmetic shift on a possibly negative value, as OpenCV
      if (h < 0)
        h += 180;
      const bool ok = h >= p.h_lo && h <= p.h_hi;
      if (ok && !in_run)
      {
        if (n_runs == 2)
          return false;
        start[n_runs] = X;
        count[n_runs] = 0;
        n_runs++;
      }
      if (ok)
        count[n_runs - 1]++;
      in_run = ok;
    }
    t.hx[diff].x = (int)(((uint32_t)start[0] & 0xffffu) | ((uint32_t)count[0] << 16));
    t.hx[diff].y = (int)(((uint32_t)start[1] & 0xffffu) | ((uint32_t)count[1] << 16));
  }
  return true;
}

__global__ void __launch_bounds__(kFastThreads) hsv_mask_flat_kernel(const uint8_t* __restrict__ bgr, unsigned long long n_px,
                                                                   const HsvDecisionTables* __restrict__ tables,
                                                                   int inverted, uint8_t* __restrict__ mask)
{
  __shared__ uint32_t s_vs[256];
  __shared__ int2 s_hx[256];
  __shared__ uint4 s_in[3 * kFastThreads];
  const unsigned long long px0 = (unsigned long long)blockIdx.x * kFastPxPerCta;
  const unsigned long long n_bytes = 3ull * n_px, byte0 = 3ull * px0;
  const uint4* src = reinterpret_cast<const uint4*>(bgr + byte0);
#pragma unroll
  for (int j = 0; j < 3; j++)
  {
    const unsigned int q = j * kFastThreads + threadIdx.x;  // 16-byte chunk of this CTA's 12 KiB
    const unsigned long long off = byte0 + 16ull * q;
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    if (off + 16 <= n_bytes)
      v = __ldg(src + q);
    else if (off < n_bytes)
    {
      // ragged end of the image: byte loads, no read past the buffer
      uint8_t tmp[16] = { 0 };
      for (int k = 0; k < 16 && off + k < n_bytes; k++)
        tmp[k] = bgr[off + k];
      memcpy(&v, tmp, 16);
    }
    s_in[q] = v;
  }
  s_vs[threadIdx.x] = __ldg(&tables->vs[threadIdx.x]);
  s_hx[threadIdx.x] = __ldg(&tables->hx[threadIdx.x]);
  __syncthreads();

  const unsigned long long first = px0 + (unsigned long long)threadIdx.x * kFastPxPerThread;
  if (first >= n_px)
    return;
  uint32_t w[12];
#pragma unroll
  for (int j = 0; j < 3; j++)
  {
    const uint4 v = s_in[3 * threadIdx.x + j];
    w[4 * j + 0] = v.x;
    w[4 * j + 1] = v.y;
    w[4 * j + 2] = v.z;
    w[4 * j + 3] = v.w;
  }
  const uint32_t inv = inverted ? 0xffffffffu : 0u;
  uint32_t out[4];
#pragma unroll
  for (int g = 0; g < 4; g++)  // 4 pixels = 12 bytes = 3 words -> 1 output word
  {
    const uint32_t a = w[3 * g], b = w[3 * g + 1], c = w[3 * g + 2];
    // pixel k of the group: bytes 3k, 3k+1, 3k+2 of the 12-byte run (little endian)
    const uint32_t m0 = decide(s_vs, s_hx, a & 0xff, (a >> 8) & 0xff, (a >> 16) & 0xff);
    const uint32_t m1 = decide(s_vs, s_hx, a >> 24, b & 0xff, (b >> 8) & 0xff);
    const uint32_t m2 = decide(s_vs, s_hx, (b >> 16) & 0xff, b >> 24, c & 0xff);
    const uint32_t m3 = decide(s_vs, s_hx, (c >> 8) & 0xff, (c >> 16) & 0xff, c >> 24);
    out[g] = (m0 | (m1 << 8) | (m2 << 16) | (m3 << 24)) ^ inv;
  }
  if (first + kFastPxPerThread <= n_px)
    *reinterpret_cast<uint4*>(mask + first) = make_uint4(out[0], out[1], out[2], out[3]);
  else
    for (unsigned long long k = 0; first + k < n_px; k++)
      mask[first + k] = (uint8_t)(out[k >> 2] >> (8 * (k & 3)));
}

// device copy of the decision tables for `p`, cached in the context; *out = NULL if `p` cannot be tabulated
int hsv_decision_tables(pcs_ctx* ctx, const pcs_hsv_params& p, const HsvDecisionTables** out)
{
  *out = nullptr;
  const bool same = ctx->hsv_tables_valid && ctx->hsv_tables_params.h_lo == p.h_lo && ctx->hsv_tables_params.h_hi == p.h_hi &&
                    ctx->hsv_tables_params.s_lo == p.s_lo && ctx->hsv_tables_params.s_hi == p.s_hi &&
                    ctx->hsv_tables_params.v_lo == p.v_lo && ctx->hsv_tables_params.v_hi == p.v_hi;
  if (!same)
  {
    static thread_local HsvDecisionTables host;  // must outlive the asynchronous copy below: synchronised before reuse
    if (ctx->hsv_tables_valid || ctx->hsv_tables_dev)
      PCS_CUDA(cudaStreamSynchronize(ctx->stream));
    if (!build_decision_tables(p, host))
      return PCS_OK;
    if (!ctx->hsv_tables_dev)
      PCS_CUDA(cudaMalloc(&ctx->hsv_tables_dev, sizeof(HsvDecisionTables)));
    PCS_CUDA(cudaMemcpyAsync(ctx->hsv_tables_dev, &host, sizeof(host), cudaMemcpyHostToDevice, ctx->stream));
    PCS_CUDA(cudaStreamSynchronize(ctx->stream));
    ctx->hsv_tables_params = p;
    ctx->hsv_tables_valid = true;
  }
  *out = (const HsvDecisionTables*)ctx->hsv_tables_dev;
  return PCS_OK;
}

int launch_hsv_mask(pcs_ctx* ctx, const uint8_t* bgr_dev, int width, int height, size_t stride,
                    const pcs_hsv_params* p, uint8_t* mask_dev, size_t mask_stride, int out_channels)
{
  HsvBounds b = { p->h_lo, p->h_hi, p->s_lo, p->s_hi, p->v_lo, p->v_hi };
  dim3 grid((unsigned)((width + kTileW - 1) / kTileW), (unsigned)((height + kTileH - 1) / kTileH));
  dim3 block(kHsvThreads);
  const int inv = p->inverted ? 1 : 0;
  const bool flat = !p->morph_open && out_channels == 1 && stride == (size_t)width * 3 && mask_stride == (size_t)width &&
                    (((uintptr_t)bgr_dev | (uintptr_t)mask_dev) & 15u) == 0;
  if (flat)
  {
    const HsvDecisionTables* tables = nullptr;
    PCS_TRY(hsv_decision_tables(ctx, *p, &tables));
    if (tables)
    {
      const unsigned long long n_px = (unsigned long long)width * height;
      hsv_mask_flat_kernel<<<(unsigned)((n_px + kFastPxPerCta - 1) / kFastPxPerCta), kFastThreads, 0, ctx->stream>>>(
          bgr_dev, n_px, tables, inv, mask_dev);
      count_launch(ctx);
      return check_launch("hsv_mask_flat_kernel");
    }
  }
  if (p->morph_open)
  {
    // tile height: 64 rows keep the halo at 8 / 64; small images get shorter tiles so that the grid covers the SMs
    const unsigned int bands = (unsigned)((width + kOpenImgTW - 1) / kOpenImgTW);
    int tile_h = kOpenImgMaxTH;
    while (tile_h > 8 && (unsigned long long)bands * ((height + tile_h - 1) / tile_h) < 2ull * (unsigned long long)ctx->sm_count)
      tile_h >>= 1;
    dim3 ogrid(bands, (unsigned)((height + tile_h - 1) / tile_h));
    if (out_channels == 1)
      hsv_open_kernel<1><<<ogrid, kOpenImgTW, 0, ctx->stream>>>(bgr_dev, width, height, stride, b, inv, mask_dev, mask_stride, tile_h);
    else
      hsv_open_kernel<3><<<ogrid, kOpenImgTW, 0, ctx->stream>>>(bgr_dev, width, height, stride, b, inv, mask_dev, mask_stride, tile_h);
    count_launch(ctx);
    return check_launch("hsv_open_kernel");
  }
  if (out_channels == 1)
    hsv_mask_kernel<1><<<grid, block, 0, ctx->stream>>>(bgr_dev, width, height, stride, b, inv, mask_dev, mask_stride);
  else
    hsv_mask_kernel<3><<<grid, block, 0, ctx->stream>>>(bgr_dev, width, height, stride, b, inv, mask_dev, mask_stride);
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
