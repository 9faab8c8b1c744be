// common.cuh -- shared declarations of libpcs_b200 (sm_100a).  Internal; the public surface is include/pcs_b200.h.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "pcs_b200.h"

namespace pcs
{
constexpr int kCopyStreams = 2;

void set_error(const char* fmt, ...);
int cuda_fail(cudaError_t e, const char* what, const char* file, int line);

#define PCS_CUDA(call)                                                                                                 \
  do                                                                                                                   \
  {                                                                                                                    \
    cudaError_t e_ = (call);                                                                                           \
    if (e_ != cudaSuccess)                                                                                             \
      return pcs::cuda_fail(e_, #call, __FILE__, __LINE__);                                                            \
  } while (0)

#define PCS_CHECK_ARG(cond, msg)                                                                                       \
  do                                                                                                                   \
  {                                                                                                                    \
    if (!(cond))                                                                                                       \
    {                                                                                                                  \
      pcs::set_error("bad argument: %s", msg);                                                                         \
      return PCS_ERR_BAD_ARG;                                                                                          \
    }                                                                                                                  \
  } while (0)

#define PCS_TRY(call)                                                                                                  \
  do                                                                                                                   \
  {                                                                                                                    \
    int s_ = (call);                                                                                                   \
    if (s_ != PCS_OK)                                                                                                  \
      return s_;                                                                                                       \
  } while (0)

// 16-byte hash slot.  key = key48 | kValidBit (0 = empty, so a plain memset clears the table);
// counts = count0 | count1 << 32, updated with one 64-bit atomic add.
struct Slot
{
  unsigned long long key;
  unsigned long long counts;
};
constexpr unsigned long long kValidBit = 1ull << 63;

struct MapView
{
  Slot* slots;
  unsigned long long mask;     // n_slots - 1 (power of two)
  unsigned int* n_used;        // distinct voxels
  unsigned int* overflow;      // set when a probe sequence wrapped
  unsigned int* live;          // live[i] = slot claimed by the i-th distinct voxel (i < *n_used): clear / push / export
                               // walk the live voxels instead of the whole table
  double inv_res;              // 1 / resolution, computed in double on the host like octomap's resolution_factor
};
}  // namespace pcs

struct pcs_map
{
  pcs::MapView view;
  double resolution;
  uint64_t capacity_voxels;
  uint64_t n_slots;
  void* block;               // one allocation: slots | live list | counters (one CUDA IPC handle exports all three)
  unsigned int* d_counters;  // [0] n_used, [1] overflow, [2..63] scratch counters, [64..79] exchange epoch flags, [127] clear
  bool is_peer;              // memory belongs to another process (opened through CUDA IPC)
};

struct pcs_ctx
{
  int device;
  cudaStream_t stream;
  bool owns_stream;
  cudaStream_t copy_stream[pcs::kCopyStreams];
  cudaEvent_t pipe_event[6];  // H2D done / kernel done / D2H done per buffer set of the host-pointer pipelines
  // grow-only device scratch (block counts, flags, staging)
  void* scratch;
  size_t scratch_bytes;
  void* staging[8];
  size_t staging_bytes[8];
  // grow-only pinned host scratch for small read-backs
  void* pinned;
  size_t pinned_bytes;
  pcs_stats stats;
  int sm_count;
  // decision tables of the flat HSV image kernel, rebuilt when the thresholds change (hsv.cu)
  void* hsv_tables_dev;
  pcs_hsv_params hsv_tables_params;
  bool hsv_tables_valid;
  void* mesh_counters;  // [0] big-triangle count, [1] finished CTAs of mesh_contacts_big_kernel; zero between calls
};

namespace pcs
{
int ctx_bind(pcs_ctx* ctx);                                   // cudaSetDevice
int ctx_scratch(pcs_ctx* ctx, size_t bytes, void** out);      // grow-only device scratch
int ctx_staging(pcs_ctx* ctx, int which, size_t bytes, void** out);
int ctx_pinned(pcs_ctx* ctx, size_t bytes, void** out);
inline void count_launch(pcs_ctx* ctx, int n = 1) { ctx->stats.kernel_launches += (uint64_t)n; }
int check_launch(const char* what);

// ---- device helpers ----------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned long long mix64(unsigned long long x)
{
  x ^= x >> 33;
  x *= 0xff51afd7ed558ccdULL;
  x ^= x >> 33;
  x *= 0xc4ceb9fe1a85ec53ULL;
  x ^= x >> 33;
  return x;
}

// 256-bit global load / store of one PointXYZRGB (sm_100: LDG.E.256 / STG.E.256)
struct Point8
{
  float x, y, z, w;
  uint32_t rgba, p0, p1, p2;
};

__device__ __forceinline__ Point8 ld_point_stream(const void* p)
{
  Point8 r;
  asm volatile("ld.global.nc.L1::no_allocate.L2::evict_first.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w), "=r"(r.rgba), "=r"(r.p0), "=r"(r.p1), "=r"(r.p2)
               : "l"(p));
  return r;
}

__device__ __forceinline__ void st_point_stream(void* p, const Point8& v)
{
#ifdef PCS_ST_EVICT_FIRST
  asm volatile("st.global.L1::no_allocate.L2::evict_first.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z),
               "f"(v.w), "r"(v.rgba), "r"(v.p0), "r"(v.p1), "r"(v.p2)
               : "memory");
#else
  asm volatile("st.global.L1::no_allocate.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z),
               "f"(v.w), "r"(v.rgba), "r"(v.p0), "r"(v.p1), "r"(v.p2)
               : "memory");
#endif
}

// octomap coordToKeyChecked (see oracle/octree.c orc_coord_to_key): k = floor(inv_res * coord) + 32768, valid iff
// 0 <= k < 65536.  F2I.FLOOR saturates (+-inf and huge values land on INT_MIN / INT_MAX, far outside the range), NaN
// converts to 0 and must be excluded by the caller's finiteness test on `coord`.
__device__ __forceinline__ bool coord_to_key(double inv_res, float coord, unsigned int& key)
{
  const int k = __double2int_rd(__dmul_rn(inv_res, (double)coord));
  key = (unsigned int)k + 32768u;
  return key < 65536u;
}

// table slot of a voxel: classic 3-prime spatial hash of the three 16-bit keys (the owner hash of the multi-GPU
// exchange is mix64, a different function, so ownership and slot position are independent)
__device__ __forceinline__ unsigned int slot_hash(unsigned long long key48)
{
  const unsigned int kx = (unsigned int)(key48 & 0xffffull), ky = (unsigned int)((key48 >> 16) & 0xffffull),
                     kz = (unsigned int)((key48 >> 32) & 0xffffull);
  unsigned int h = (kx * 73856093u) ^ (ky * 19349663u) ^ (kz * 83492791u);
  return h ^ (h >> 15);
}

__device__ __forceinline__ unsigned long long ld_relaxed_u64(const unsigned long long* p)
{
  unsigned long long v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}

// add (count0 | count1 << 32) to the voxel `key48`, inserting it if needed.  Linear probing, load-before-CAS.
__device__ __forceinline__ void map_add(const MapView& m, unsigned long long key48, unsigned long long add)
{
  const unsigned long long want = key48 | kValidBit;
  unsigned long long slot = slot_hash(key48) & m.mask;
  // the probe cap keeps a (mis-sized) full table from turning every insert into a scan of the whole table
#pragma unroll 1
  for (unsigned int probe = 0; probe < 65536u; ++probe)
  {
    Slot* s = m.slots + slot;
    unsigned long long k = ld_relaxed_u64(&s->key);
    if (k == 0ull)
    {
      k = atomicCAS(&s->key, 0ull, want);
      if (k == 0ull)
      {
        m.live[atomicAdd(m.n_used, 1u)] = (unsigned int)slot;
        k = want;
      }
    }
    if (k == want)
    {
      atomicAdd(&s->counts, add);
      return;
    }
    slot = (slot + 1) & m.mask;
  }
  *m.overflow = 1u;
}

// Insert for kernels whose threads mostly CLAIM new slots (merging records into an empty table): the claim is reported
// instead of being appended to the live list, and live_append_warp() appends all claims of a warp with ONE atomic on
// n_used -- a returning atomic per voxel on that single address serialises (~1 ns each) and dominated those kernels.
__device__ __forceinline__ bool map_add_claim(const MapView& m, unsigned long long key48, unsigned long long add,
                                              unsigned int* slot_out)
{
  const unsigned long long want = key48 | kValidBit;
  unsigned long long slot = slot_hash(key48) & m.mask;
#pragma unroll 1
  for (unsigned int probe = 0; probe < 65536u; ++probe)
  {
    Slot* s = m.slots + slot;
    unsigned long long k = ld_relaxed_u64(&s->key);
    bool claimed = false;
    if (k == 0ull)
    {
      k = atomicCAS(&s->key, 0ull, want);
      claimed = (k == 0ull);
      if (claimed)
        k = want;
    }
    if (k == want)
    {
      atomicAdd(&s->counts, add);
      *slot_out = (unsigned int)slot;
      return claimed;
    }
    slot = (slot + 1) & m.mask;
  }
  *m.overflow = 1u;
  return false;
}

// all 32 lanes must call this (converged); `claimed` lanes get consecutive live-list positions from one atomic
__device__ __forceinline__ void live_append_warp(const MapView& m, bool claimed, unsigned int slot)
{
  const unsigned int grp = __ballot_sync(0xffffffffu, claimed);
  if (!grp)
    return;
  const unsigned int lane = threadIdx.x & 31u;
  const int leader = __ffs(grp) - 1;
  unsigned int base = 0;
  if ((int)lane == leader)
    base = atomicAdd(m.n_used, (unsigned int)__popc(grp));
  base = __shfl_sync(0xffffffffu, base, leader);
  if (claimed)
    m.live[base + __popc(grp & ((1u << lane) - 1u))] = slot;
}

// warp-aggregated insert: lanes of `valid` points with equal keys elect a leader that adds the group's counts once
__device__ __forceinline__ void map_add_warp(const MapView& m, bool valid, unsigned long long key48, bool cls1)
{
  const unsigned active = __ballot_sync(0xffffffffu, valid);
  const unsigned ones = __ballot_sync(0xffffffffu, valid && cls1);
  if (valid)
  {
    const unsigned peers = __match_any_sync(active, key48);
    const unsigned lane = threadIdx.x & 31u;
    if ((unsigned)(__ffs(peers) - 1) == lane)
    {
      const unsigned c1 = __popc(peers & ones);
      const unsigned c0 = __popc(peers) - c1;
      map_add(m, key48, (unsigned long long)c0 | ((unsigned long long)c1 << 32));
    }
  }
}

// Split form of map_add_warp for kernels that hold several points per thread: _begin elects the leaders and issues the
// load of each leader's home slot, _end completes the insert.  Calling _begin for all points before the first _end keeps
// one table lookup (an L2 round trip) in flight per point instead of serialising them.
struct MapProbe
{
  unsigned long long want, add, slot, k;
  bool leader;
};

__device__ __forceinline__ MapProbe map_add_warp_begin(const MapView& m, bool valid, unsigned long long key48, bool cls1)
{
  MapProbe pr;
  pr.leader = false;
  pr.want = pr.add = pr.slot = pr.k = 0ull;
  const unsigned active = __ballot_sync(0xffffffffu, valid);
  const unsigned ones = __ballot_sync(0xffffffffu, valid && cls1);
  if (valid)
  {
    const unsigned peers = __match_any_sync(active, key48);
    const unsigned lane = threadIdx.x & 31u;
    if ((unsigned)(__ffs(peers) - 1) == lane)
    {
      const unsigned c1 = __popc(peers & ones);
      const unsigned c0 = __popc(peers) - c1;
      pr.leader = true;
      pr.add = (unsigned long long)c0 | ((unsigned long long)c1 << 32);
      pr.want = key48 | kValidBit;
      pr.slot = slot_hash(key48) & m.mask;
      pr.k = ld_relaxed_u64(&m.slots[pr.slot].key);
    }
  }
  return pr;
}

__device__ __forceinline__ void map_add_warp_end(const MapView& m, const MapProbe& pr)
{
  if (!pr.leader)
    return;
  unsigned long long slot = pr.slot, k = pr.k;
#pragma unroll 1
  for (unsigned int probe = 0; probe < 65536u; ++probe)
  {
    Slot* s = m.slots + slot;
    if (probe)
      k = ld_relaxed_u64(&s->key);
    if (k == 0ull)
    {
      k = atomicCAS(&s->key, 0ull, pr.want);
      if (k == 0ull)
      {
        m.live[atomicAdd(m.n_used, 1u)] = (unsigned int)slot;
        k = pr.want;
      }
    }
    if (k == pr.want)
    {
      atomicAdd(&s->counts, pr.add);
      return;
    }
    slot = (slot + 1) & m.mask;
  }
  *m.overflow = 1u;
}

__device__ __forceinline__ bool color_pass(int r, int g, int b, int lower, int upper, int neg)
{
  if (!neg)
    return (r >= lower && r <= upper) && (g >= lower && g <= upper) && (b >= lower && b <= upper);
  return (r <= lower || r >= upper) && (g <= lower || g >= upper) && (b <= lower || b >= upper);
}

__device__ __forceinline__ bool finite3(float x, float y, float z)
{
  // all three exponents below 0xff
  return ((__float_as_uint(x) & 0x7f800000u) != 0x7f800000u) && ((__float_as_uint(y) & 0x7f800000u) != 0x7f800000u) &&
         ((__float_as_uint(z) & 0x7f800000u) != 0x7f800000u);
}

// pcl::transformPointCloud arithmetic: ((m0 x + m1 y) + m2 z) + m3 in float, never contracted to FMA
__device__ __forceinline__ float xform_row(float m0, float m1, float m2, float m3, float x, float y, float z)
{
  return __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(m0, x), __fmul_rn(m1, y)), __fmul_rn(m2, z)), m3);
}
}  // namespace pcs
