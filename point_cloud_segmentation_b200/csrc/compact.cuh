// compact.cuh -- order-preserving stream compaction in three launches (count, scan of block counts, scatter).
// Used where the reference's output order is part of the contract: colorPassthrough keeps input order
// (octomap_mesh_masking.cpp:71-94) and maskMesh emits surviving triangles in input order (.cpp:250-272).
// Inside a block the order comes from warp ballots; across blocks from an exclusive scan of per-block counts kept in
// context scratch memory.
#pragma once

#include "common.cuh"

namespace pcs
{
constexpr int kCompactThreads = 256;
constexpr int kCompactRounds = 8;
constexpr int kCompactItems = kCompactThreads * kCompactRounds;  // items per block

template <class Pred>
__global__ void __launch_bounds__(kCompactThreads) compact_count_kernel(Pred pred, unsigned long long n,
                                                                       unsigned int* __restrict__ block_counts)
{
  __shared__ unsigned int s_warp[kCompactThreads / 32];
  const unsigned long long base = (unsigned long long)blockIdx.x * kCompactItems;
  unsigned int mine = 0;
#pragma unroll
  for (int r = 0; r < kCompactRounds; r++)
  {
    const unsigned long long i = base + (unsigned long long)r * kCompactThreads + threadIdx.x;
    mine += (i < n && pred(i)) ? 1u : 0u;
  }
  for (int o = 16; o > 0; o >>= 1)
    mine += __shfl_xor_sync(0xffffffffu, mine, o);
  if ((threadIdx.x & 31) == 0)
    s_warp[threadIdx.x >> 5] = mine;
  __syncthreads();
  if (threadIdx.x == 0)
  {
    unsigned int t = 0;
    for (int w = 0; w < kCompactThreads / 32; w++)
      t += s_warp[w];
    block_counts[blockIdx.x] = t;
  }
}

// single block: exclusive scan of n_blocks counts into 64-bit offsets; offsets[n_blocks] = total
static __global__ void __launch_bounds__(1024) compact_scan_kernel(const unsigned int* __restrict__ counts,
                                                            unsigned long long* __restrict__ offsets,
                                                            unsigned int n_blocks)
{
  __shared__ unsigned long long s_warp[32];
  __shared__ unsigned long long s_carry;
  if (threadIdx.x == 0)
    s_carry = 0;
  __syncthreads();
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (unsigned int base = 0; base < n_blocks; base += 1024)
  {
    const unsigned int i = base + threadIdx.x;
    const unsigned long long v = (i < n_blocks) ? counts[i] : 0ull;
    unsigned long long incl = v;
    for (int o = 1; o < 32; o <<= 1)
    {
      const unsigned long long t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= (unsigned)o)
        incl += t;
    }
    if (lane == 31)
      s_warp[warp] = incl;
    __syncthreads();
    if (warp == 0)
    {
      unsigned long long w = s_warp[lane];
      unsigned long long wi = w;
      for (int o = 1; o < 32; o <<= 1)
      {
        const unsigned long long t = __shfl_up_sync(0xffffffffu, wi, o);
        if (lane >= (unsigned)o)
          wi += t;
      }
      s_warp[lane] = wi - w;  // exclusive prefix of warp totals
    }
    __syncthreads();
    const unsigned long long carry = s_carry;
    if (i < n_blocks)
      offsets[i] = carry + s_warp[warp] + incl - v;
    __syncthreads();
    if (threadIdx.x == 1023)
      s_carry = carry + s_warp[warp] + incl;
    __syncthreads();
  }
  if (threadIdx.x == 0)
    offsets[n_blocks] = s_carry;
}

template <class Pred, class Emit>
__global__ void __launch_bounds__(kCompactThreads) compact_scatter_kernel(Pred pred, Emit emit, unsigned long long n,
                                                                         const unsigned long long* __restrict__ offsets)
{
  __shared__ unsigned int s_warp[kCompactThreads / 32];
  const unsigned long long base = (unsigned long long)blockIdx.x * kCompactItems;
  unsigned long long running = offsets[blockIdx.x];
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int r = 0; r < kCompactRounds; r++)
  {
    const unsigned long long i = base + (unsigned long long)r * kCompactThreads + threadIdx.x;
    const bool keep = (i < n) && pred(i);
    const unsigned ballot = __ballot_sync(0xffffffffu, keep);
    if (lane == 0)
      s_warp[warp] = __popc(ballot);
    __syncthreads();
    unsigned int before = 0, total = 0;
#pragma unroll
    for (int w = 0; w < kCompactThreads / 32; w++)
    {
      const unsigned int c = s_warp[w];
      before += (w < (int)warp) ? c : 0u;
      total += c;
    }
    if (keep)
      emit(i, running + before + __popc(ballot & ((1u << lane) - 1u)));
    running += total;
    __syncthreads();
  }
}

inline unsigned int compact_blocks(unsigned long long n)
{
  return (unsigned int)((n + kCompactItems - 1) / kCompactItems);
}

// scratch layout: [n_blocks] u32 counts (padded to 8 B) then [n_blocks + 1] u64 offsets
inline size_t compact_scratch_bytes(unsigned long long n)
{
  const size_t nb = compact_blocks(n);
  return ((nb * sizeof(unsigned int) + 15) & ~(size_t)15) + (nb + 1) * sizeof(unsigned long long);
}

// Count + scan only: per-block exclusive offsets for a caller that brings its own scatter kernel (same block shape:
// kCompactThreads threads, kCompactRounds rounds, item = base + round * kCompactThreads + thread).
template <class Pred>
int compact_offsets(pcs_ctx* ctx, Pred pred, unsigned long long n, void* scratch, unsigned long long** offsets_out)
{
  const unsigned int nb = compact_blocks(n);
  unsigned int* counts = reinterpret_cast<unsigned int*>(scratch);
  unsigned long long* offsets = reinterpret_cast<unsigned long long*>(
      reinterpret_cast<char*>(scratch) + (((size_t)nb * sizeof(unsigned int) + 15) & ~(size_t)15));
  if (nb > 0)
  {
    compact_count_kernel<<<nb, kCompactThreads, 0, ctx->stream>>>(pred, n, counts);
    count_launch(ctx);
  }
  compact_scan_kernel<<<1, 1024, 0, ctx->stream>>>(counts, offsets, nb);
  count_launch(ctx);
  *offsets_out = offsets;
  return check_launch("compact_offsets");
}

// Enqueues the three launches.  `scratch` must hold compact_scratch_bytes(n).  The total lands at offsets[n_blocks];
// its device address is returned through d_total.
template <class Pred, class Emit>
int compact_run(pcs_ctx* ctx, Pred pred, Emit emit, unsigned long long n, void* scratch,
                const unsigned long long** d_total)
{
  const unsigned int nb = compact_blocks(n);
  unsigned long long* offsets = nullptr;
  PCS_TRY(compact_offsets(ctx, pred, n, scratch, &offsets));
  if (nb > 0)
  {
    compact_scatter_kernel<<<nb, kCompactThreads, 0, ctx->stream>>>(pred, emit, n, offsets);
    count_launch(ctx);
  }
  *d_total = offsets + nb;
  return check_launch("compact");
}
}  // namespace pcs
