// annotate_open.cu -- the fused annotate(+integrate) pass with the detector's 5x5 opening, still ONE launch per batch.
//
// Replaces the same call chain as annotate.cu (cloudToImage -> detector -> imageToCloud -> transformPointCloud ->
// colorPassthrough -> octree insert), with the detector being HSV threshold + MORPH_OPEN 5x5 (semantics of
// pcs_detection/src_python/pcs_detection/models/histogram_backprojection.py:40-41, cv2.morphologyEx on a {0,255} mask).
// An opened mask pixel depends on the 9x9 neighbourhood of thresholded pixels, so the kernel works on the organized grid:
//
//   * a CTA owns a band of 256 columns and a section of ~180 rows of one frame and marches down the rows, 4 rows (a chunk)
//     at a time.  A PRODUCER warp brings each row segment (256 + 4 + 4 halo points, 8.25 KiB contiguous) into a 12-row
//     shared-memory ring with one TMA bulk copy (cp.async.bulk ... mbarrier::complete_tx, L2::evict_first), chunk after
//     chunk, as soon as the consumers have released the ring chunk (`empty` mbarriers); no registers are spent on staging
//     and every point is read from HBM once (+ halo: 4 rows per section end, 4 columns per band side; no BGR plane, no mask
//     image, no second pass over the cloud);
//   * 16 CONSUMER warps (two groups of 8, each group takes 2 of the 4 rows of a chunk) wait on the chunk's `full` mbarrier,
//     threshold their rows with host-built decision tables (hsv.cuh) and bit-pack the result with warp ballots (32 pixels per
//     word).  A word column (32 columns) thresholds its 4 + 4 neighbouring pixels again instead of fetching them from the
//     neighbouring warps, so the word columns run independently of each other: there is no block-wide barrier per chunk,
//     only a named barrier between the two warps of a column;
//   * the opening is separable and runs on the packed words: a 5-wide min is 3 shift/AND steps on a 64-bit window, a 5-tall
//     min is 4 ANDs of rows held by neighbouring lanes (shuffles), dilation likewise with ORs; group 0's warp does it for
//     the column and publishes the opened rows (the instruction count of this step does not depend on the lanes in use);
//   * four rows behind the threshold front the mask is final and the points -- still in the ring -- are finished exactly as
//     in annotate.cu (label, transform, annotated cloud + mask stores, warp-aggregated voxel update; the table lookups of a
//     thread's rows are issued back to back).
//
// Border rule of cv2 (BORDER_CONSTANT with morphologyDefaultBorderValue): pixels outside the image do not take part in
// either the erosion or the dilation = thresholded pixels outside the image count as 1, eroded pixels outside as 0.
//
// Algorithmic bytes per point: 65 (same as annotate.cu) + halo re-reads (8 rows per 180-row section, 8 columns per 256).
#include <stdlib.h>

#include "annotate.cuh"

namespace pcs
{
#ifndef PCS_OPEN_TW
#define PCS_OPEN_TW 256
#endif
#ifndef PCS_OPEN_CHUNK
#define PCS_OPEN_CHUNK 4
#endif
#ifndef PCS_OPEN_GROUPS
#define PCS_OPEN_GROUPS 2
#endif
constexpr int kOpenTW = PCS_OPEN_TW;                          // band width in columns
constexpr int kOpenWords = kOpenTW / 32;                      // packed words per band row
constexpr int kOpenChunk = PCS_OPEN_CHUNK;                    // rows per TMA chunk / per iteration (>= 4, multiple of 4)
constexpr int kOpenGroups = PCS_OPEN_GROUPS;                  // thread groups per CTA, each takes chunk / groups rows
constexpr int kOpenConsumers = kOpenTW * kOpenGroups;          // threads that threshold / open / finish
constexpr int kOpenThreads = kOpenConsumers + 32;              // + one producer warp that only issues the TMA copies
constexpr int kOpenRingRows = 3 * kOpenChunk;                 // rows being finished, thresholded, and in flight
constexpr int kOpenRowPts = kOpenTW + 8;                      // 4 halo columns on each side
constexpr int kOpenRowBytes = kOpenRowPts * 32;
constexpr int kOpenRawRows = (3 * kOpenChunk > 2 * kOpenChunk + 8) ? 3 * kOpenChunk : 2 * kOpenChunk + 8;
static_assert(kOpenChunk % 4 == 0 && kOpenChunk >= 4 && kOpenChunk + 8 <= 32, "chunk must be 4, 8, .. 24");
static_assert(kOpenChunk % kOpenGroups == 0, "groups must divide the chunk");
static_assert(8 * kOpenChunk <= 32, "halo pass: 8 columns x chunk rows must fit a warp");
static_assert(kOpenWords <= 15, "one named barrier per word column");

struct OpenSmem
{
  unsigned char ring[kOpenRingRows][kOpenRowBytes];
  // packed threshold rows, private to each 32-column word column: [0] the 32 own columns, [1] the 4 + 4 halo columns
  unsigned int raw[kOpenWords][kOpenRawRows][2];
  uint32_t vs[256];
  int2 hx[256];
  unsigned int fin[kOpenWords][kOpenChunk];  // opened rows of the current iteration, published by group 0's warp
  unsigned long long full[3];               // ring chunk landed (TMA transaction bytes)
  unsigned long long empty[3];              // ring chunk finished by every consumer warp
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(unsigned long long* bar, unsigned int count)
{
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}

__device__ __forceinline__ void mbar_arrive_expect_tx(unsigned long long* bar, unsigned int bytes)
{
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}

__device__ __forceinline__ void mbar_arrive(unsigned long long* bar)
{
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

__device__ __forceinline__ bool mbar_try_wait(unsigned long long* bar, unsigned int parity)
{
  uint32_t ok;
  asm volatile("{\n"
               "  .reg .pred p;\n"
               "  mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
               "  selp.u32 %0, 1, 0, p;\n"
               "}\n"
               : "=r"(ok)
               : "r"(smem_u32(bar)), "r"(parity)
               : "memory");
  return ok != 0;
}

// global -> shared bulk copy through the TMA unit; completion is counted in bytes on `bar`
__device__ __forceinline__ void tma_load_1d(void* dst_smem, const void* src_gmem, unsigned int bytes, unsigned long long* bar,
                                            unsigned long long policy)
{
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(
                   smem_u32(dst_smem)),
               "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)), "l"(policy)
               : "memory");
}

template <bool HAS_MAP, bool WRITE_ANN, bool WRITE_MASK>
__global__ void __launch_bounds__(kOpenThreads, 2)
    annotate_open_kernel(const AnnotateArgs a, const unsigned int bands, const unsigned int sections, const unsigned int section_rows)
{
  extern __shared__ __align__(128) unsigned char open_smem[];
  OpenSmem& S = *reinterpret_cast<OpenSmem*>(open_smem);
  constexpr int CH = kOpenChunk;
  constexpr int kRowsPerThread = CH / kOpenGroups;
  constexpr unsigned int kConsumerWarps = kOpenConsumers / 32;
  const unsigned int lane = threadIdx.x & 31u;
  const unsigned int wid = threadIdx.x >> 5;                        // warp in the CTA; the last one is the producer

  unsigned int blk = blockIdx.x;
  const unsigned int band = blk % bands;
  blk /= bands;
  const unsigned int section = blk % sections;
  const unsigned int frame = blk / sections;

  const int W = (int)a.width, H = (int)a.height;
  const int x0 = (int)(band * kOpenTW);
  const int r0 = (int)(section * section_rows);
  const int r1 = min(r0 + (int)section_rows, H);
  if (r0 >= r1)
    return;  // uniform: the last section of a short frame can be empty
  // extended row e (0-based from r0 - 4) <-> image row y = r0 - 4 + e; chunk k holds e = CH k .. CH k + CH - 1
  const int load_end = min(H, r1 + 4);                            // rows [max(0, r0 - 4), load_end) are brought in
  const int n_chunks = (r1 - r0 + 8 + CH - 1) / CH;

  if (threadIdx.x == 0)
  {
    for (int i = 0; i < 3; i++)
    {
      mbar_init(&S.full[i], 1u);
      mbar_init(&S.empty[i], kConsumerWarps);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  for (unsigned int i = threadIdx.x; i < 256u; i += kOpenThreads)
  {
    S.vs[i] = __ldg(&a.hsv_tables->vs[i]);
    S.hx[i] = __ldg(&a.hsv_tables->hx[i]);
  }
  __syncthreads();  // the only block-wide barrier: from here on the warps synchronise through mbarriers / pair barriers

  if (wid == kConsumerWarps)
  {
    // ---- producer warp: one lane streams the row segments of chunk after chunk into the ring -----------------------------
    if (lane == 0)
    {
      const int cs = max(x0 - 4, 0), ce = min(x0 + kOpenTW + 4, W);  // columns of a row segment
      const unsigned int row_bytes = (unsigned int)(ce - cs) * 32u;
      const unsigned int dst_off = (unsigned int)(cs - (x0 - 4)) * 32u;  // column c lives at (c - (x0 - 4)) * 32 in a ring row
      const Point8* frame_pts = a.cloud + (size_t)frame * a.pts_per_frame;
      unsigned long long policy;
      asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(policy));
      for (int k = 0; k < n_chunks; k++)
      {
        // the slots of chunk k held chunk k - 3: wait until every consumer warp has finished its points
        if (k >= 3)
          while (!mbar_try_wait(&S.empty[k % 3], (unsigned int)(k / 3 - 1) & 1u))
          {
          }
        unsigned long long* bar = &S.full[k % 3];
        unsigned int bytes = 0;
#pragma unroll
        for (int i = 0; i < CH; i++)
        {
          const int y = r0 - 4 + CH * k + i;
          bytes += (y >= 0 && y < load_end) ? row_bytes : 0u;
        }
        mbar_arrive_expect_tx(bar, bytes);  // zero bytes: the phase completes with this arrival
#pragma unroll
        for (int i = 0; i < CH; i++)
        {
          const int e = CH * k + i;
          const int y = r0 - 4 + e;
          if (y >= 0 && y < load_end)
            tma_load_1d(&S.ring[e % kOpenRingRows][dst_off], frame_pts + (size_t)y * W + cs, row_bytes, bar, policy);
        }
      }
    }
    return;
  }

  // ---- consumer warps ------------------------------------------------------------------------------------------------------
  const unsigned int tid = threadIdx.x % kOpenTW;                  // column inside the band
  const unsigned int warp = tid >> 5;                              // word column of this warp
  const unsigned int grp = threadIdx.x / kOpenTW;                  // which rows of a chunk this thread takes
  const int xw = x0 + 32 * (int)warp;                              // first column of the word column

  Pose P;
  const bool has_pose = (a.poses != nullptr);
  if (has_pose && (HAS_MAP || WRITE_ANN))
    load_pose(a.poses, frame, P);

  // validity of the 64 columns [xw - 16, xw + 48) this warp's erosion window covers (eroded pixels outside the image are 0)
  unsigned long long col_valid = ~0ull;
  {
    const int lo = xw - 16;
    if (lo < 0)
      col_valid &= ~0ull << (-lo);
    const int n = W - lo;  // columns lo .. W-1 are inside
    if (n < 64)
      col_valid &= (n <= 0) ? 0ull : ((1ull << n) - 1ull);
  }
  const int col = x0 + (int)tid;
  const bool col_ok = col < W;
  const unsigned int pt_off = (tid + 4u) * 32u;  // this thread's column inside a ring row
  // halo pass of A (group 0's warp, all rows of the chunk): lane = 8 * row + halo column, the 4 columns left and the 4
  // right of the word column
  const int h_row = (int)(lane >> 3), hc = (int)(lane & 7u);
  const int h_col = hc < 4 ? xw - 4 + hc : xw + 32 + (hc - 4);
  const bool h_ok = h_row < CH && h_col >= 0 && h_col < W;
  const unsigned int h_off = (unsigned int)(h_col - (x0 - 4)) * 32u + 16u;
  const unsigned int pair_bar = 1u + warp, pair_threads = 32u * kOpenGroups;

  for (int k = 0; k < n_chunks; k++)
  {
    // ---- A(k): threshold this thread's rows of the chunk that has just landed, pack them ----------------------------------
    while (!mbar_try_wait(&S.full[k % 3], (unsigned int)(k / 3) & 1u))
    {
    }
    const int e0 = CH * k;  // first extended row of the chunk
#pragma unroll
    for (int j = 0; j < kRowsPerThread; j++)
    {
      const int e = e0 + (int)grp * kRowsPerThread + j;
      const int y = r0 - 4 + e;
      bool bit = true;  // outside the image: does not constrain the erosion
      if (col_ok && y >= 0 && y < load_end)
      {
        const unsigned int rgba = *reinterpret_cast<const unsigned int*>(&S.ring[e % kOpenRingRows][pt_off + 16u]);
        bit = decide(S.vs, S.hx, rgba & 0xff, (rgba >> 8) & 0xff, (rgba >> 16) & 0xff) != 0u;
      }
      const unsigned int word = __ballot_sync(0xffffffffu, bit);
      if (lane == 0)
        S.raw[warp][e % kOpenRawRows][0] = word;
    }
    if (grp == 0)
    {
      // the neighbours' 4 + 4 columns are thresholded again here rather than fetched from the neighbouring warps: that
      // keeps the word columns independent of each other (no block-wide barrier per chunk)
      const int e = e0 + h_row;
      const int y = r0 - 4 + e;
      bool bit = true;
      if (h_ok && y >= 0 && y < load_end)
      {
        const unsigned int rgba = *reinterpret_cast<const unsigned int*>(&S.ring[e % kOpenRingRows][h_off]);
        bit = decide(S.vs, S.hx, rgba & 0xff, (rgba >> 8) & 0xff, (rgba >> 16) & 0xff) != 0u;
      }
      const unsigned int hb = __ballot_sync(0xffffffffu, bit);
      if (lane < (unsigned int)CH)
        S.raw[warp][(e0 + (int)lane) % kOpenRawRows][1] = (hb >> (8u * lane)) & 0xffu;
    }
    // rows that can be finished now: extended rows f0 .. f0 + CH - 1 (thresholded rows up to e0 + CH - 1 are known)
    const int f0 = e0 - 4;
    const bool finishing = f0 + CH - 1 >= 4;  // else still inside the upper halo (uniform)

    // ---- B(k): open this word column's rows on the packed words --------------------------------------------------------------
    // Group 0's warp does it for all groups of the column (the instruction count of B does not depend on how many lanes
    // are busy); the column's warps meet at a named barrier before (threshold rows written) and after (opened rows
    // published).
    if (kOpenGroups > 1)
      asm volatile("bar.sync %0, %1;" ::"r"(pair_bar), "r"(pair_threads) : "memory");
    else
      __syncwarp();
    unsigned int fin = 0;
    if (finishing)
    {
      if (grp == 0)
      {
        // lane l < CH + 8: horizontal 5-min of thresholded row f0 - 4 + l over the 64-bit window [xw - 16, xw + 48)
        const int er = max(f0 - 4 + (int)min(lane, (unsigned int)(CH + 7)), 0);
        const uint2 rw = *reinterpret_cast<const uint2*>(S.raw[warp][er % kOpenRawRows]);
        const unsigned long long win = ((unsigned long long)(rw.y & 0xfu) << 12) | ((unsigned long long)rw.x << 16) |
                                       ((unsigned long long)(rw.y >> 4) << 48);
        // 5-wide min in three shift/AND steps: pairs, triples (bits i .. i+2), then triples at i-2 and i
        const unsigned long long w2 = win & (win >> 1), w3 = w2 & (win >> 2);
        const unsigned long long he = w3 & (w3 << 2);
        // lane l < CH + 4: eroded row f0 - 2 + l = AND of he rows l .. l+4
        unsigned long long er5 = he;
#pragma unroll
        for (int d = 1; d <= 4; d++)
          er5 &= __shfl_down_sync(0xffffffffu, he, d);
        const int y_e = r0 - 4 + f0 - 2 + (int)lane;
        er5 = (y_e >= 0 && y_e < H) ? (er5 & col_valid) : 0ull;
        const unsigned long long d2 = er5 | (er5 >> 1), d3 = d2 | (er5 >> 2);
        const unsigned long long hd = d3 | (d3 << 2);
        // lane l < CH: opened row f0 + l = OR of hd rows l .. l+4
        unsigned long long op = hd;
#pragma unroll
        for (int d = 1; d <= 4; d++)
          op |= __shfl_down_sync(0xffffffffu, hd, d);
        fin = (unsigned int)(op >> 16);
        if (kOpenGroups > 1 && lane < (unsigned int)CH)
          S.fin[warp][lane] = fin;
      }
      if (kOpenGroups > 1)
      {
        asm volatile("bar.sync %0, %1;" ::"r"(pair_bar), "r"(pair_threads) : "memory");
        if (grp != 0)
          fin = S.fin[warp][min(lane, (unsigned int)(CH - 1))];
      }

      // ---- C(k): finish the points of those rows -------------------------------------------------------------------------------
      // outputs of this thread's rows first, with their voxel-table lookups issued back to back; the inserts follow
      MapProbe probe[kRowsPerThread];
#pragma unroll
      for (int j = 0; j < kRowsPerThread; j++)
      {
        const int f = (int)grp * kRowsPerThread + j;
        const unsigned int row_bits = __shfl_sync(0xffffffffu, fin, f);
        const int e = f0 + f;
        const int y = r0 - 4 + e;
        const bool in_frame = y >= r0 && y < r1 && col_ok;  // rows outside the section belong to its neighbours (halo here)
        const uint4* src = reinterpret_cast<const uint4*>(&S.ring[(e + kOpenRingRows) % kOpenRingRows][pt_off]);
        const uint4 lo = src[0], hi = src[1];
        Point8 p;
        p.x = __uint_as_float(lo.x);
        p.y = __uint_as_float(lo.y);
        p.z = __uint_as_float(lo.z);
        p.w = __uint_as_float(lo.w);
        p.rgba = hi.x;
        p.p0 = hi.y;
        p.p1 = hi.z;
        p.p2 = hi.w;
        const bool det = (((row_bits >> lane) & 1u) != 0u) != (a.inverted != 0);
        bool valid, cls1;
        unsigned long long key;
        const long long idx = (long long)frame * a.pts_per_frame + (long long)y * W + col;
        emit_point<HAS_MAP, WRITE_ANN, WRITE_MASK>(a, P, has_pose, p, (unsigned long long)idx, in_frame, det, valid, key, cls1);
        if (HAS_MAP)
          probe[j] = map_add_warp_begin(a.map, valid, key, cls1);
      }
      if (HAS_MAP)
      {
#pragma unroll
        for (int j = 0; j < kRowsPerThread; j++)
          map_add_warp_end(a.map, probe[j]);
      }
    }
    // this warp is done with the ring rows of chunk k - 1 (thresholded at iteration k - 1, finished just now)
    if (k >= 1)
    {
      __syncwarp();
      if (lane == 0)
        mbar_arrive(&S.empty[(k - 1) % 3]);
    }
  }
}

int launch_annotate_open(pcs_ctx* ctx, const AnnotateArgs& a, unsigned int n_frames, bool hm, bool wa, bool wm)
{
  const unsigned int W = a.width, H = a.height;
  const unsigned int bands = (W + kOpenTW - 1) / kOpenTW;
  // Section height: ~180 rows keeps the vertical halo re-read at 8 / 180; small batches get shorter sections so that the
  // grid still fills the machine (a single 640x480 frame: 16-row sections, 90 CTAs).
  unsigned int s_min = (H + 90) / 180, s_max = H / 16;
  if (s_min < 1)
    s_min = 1;
  if (s_max < 1)
    s_max = 1;
  const unsigned long long cols = (unsigned long long)n_frames * bands;
  unsigned long long want = ((unsigned long long)ctx->sm_count * 4ull + cols - 1) / cols;
  unsigned int sections = (unsigned int)(want < s_min ? s_min : (want > s_max ? s_max : want));
  unsigned int section_rows = (H + sections - 1) / sections;
  if (const char* env = getenv("PCS_OPEN_SECTION_ROWS"))  // tuning / test aid: force the section height
    if (atoi(env) > 0)
      section_rows = (unsigned int)atoi(env);
  section_rows = (section_rows + kOpenChunk - 1) / kOpenChunk * kOpenChunk;
  sections = (H + section_rows - 1) / section_rows;
  const unsigned long long grid = cols * sections;
  if (grid > 0x7fffffffull)
  {
    set_error("batch too large for one launch: %llu CTAs; split the batch", grid);
    return PCS_ERR_BAD_ARG;
  }
  const size_t smem = sizeof(OpenSmem) + 128;
#define PCS_OPEN_CASE(HM, WA, WM)                                                                                      \
  if (hm == HM && wa == WA && wm == WM)                                                                                \
  {                                                                                                                    \
    static bool configured[64] = {};                                                                                   \
    if (ctx->device < 64 && !configured[ctx->device])                                                                  \
    {                                                                                                                  \
      PCS_CUDA(cudaFuncSetAttribute(annotate_open_kernel<HM, WA, WM>, cudaFuncAttributeMaxDynamicSharedMemorySize,     \
                                    (int)smem));                                                                       \
      configured[ctx->device] = true;                                                                                  \
    }                                                                                                                  \
    annotate_open_kernel<HM, WA, WM><<<(unsigned int)grid, kOpenThreads, smem, ctx->stream>>>(a, bands, sections, section_rows); \
  }
  PCS_OPEN_CASE(true, true, true)
  PCS_OPEN_CASE(true, true, false)
  PCS_OPEN_CASE(true, false, true)
  PCS_OPEN_CASE(true, false, false)
  PCS_OPEN_CASE(false, true, true)
  PCS_OPEN_CASE(false, true, false)
  PCS_OPEN_CASE(false, false, true)
#undef PCS_OPEN_CASE
  count_launch(ctx);
  return check_launch("annotate_open_kernel");
}
}  // namespace pcs
