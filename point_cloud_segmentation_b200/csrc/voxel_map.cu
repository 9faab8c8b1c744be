// voxel_map.cu -- stage 3 state: a device hash of 48-bit voxel keys standing in for the octomap::OcTree that
// OctomapMeshMask::setOctree builds through tesseract_geometry::Octree
// (pcs_scan_integration/src/pcs_scan_integration/octomap_mesh_masking.cpp:105-116).
//
// The path inserts hits only, so a leaf's log-odds is a pure function of its hit count (octomap: +logodds(0.7) per
// hit, clamped at logodds(0.971), early return once clamped) and the pointer tree reduces to a count per key.
// Export re-creates the octree's view: records in Morton (child-index DFS) order with the float log-odds.
#include <cub/device/device_radix_sort.cuh>
#include <cuda.h>

#include <math.h>
#include <string.h>
#include <new>

#include "common.cuh"

namespace pcs
{
struct Record
{
  unsigned long long key48;
  unsigned long long counts;  // count0 | count1 << 32
};

__device__ __forceinline__ unsigned long long part1by2(unsigned long long x)
{
  // spread the low 16 bits so that bit i lands at bit 3i
  x &= 0xffffull;
  x = (x | (x << 16)) & 0x000000ff0000ffull;
  x = (x | (x << 8)) & 0x00f00f00f00f00full;
  x = (x | (x << 4)) & 0x0c30c30c30c30c3ull;
  x = (x | (x << 2)) & 0x249249249249249ull;
  return x;
}

__device__ __forceinline__ unsigned long long morton48(unsigned long long key48)
{
  // octomap child index: bit0 -> x, bit1 -> y, bit2 -> z at every level
  return part1by2(key48) | (part1by2(key48 >> 16) << 1) | (part1by2(key48 >> 32) << 2);
}

// gather the used slots (arbitrary order) and their Morton codes
__global__ void map_gather_kernel(MapView m, Record* __restrict__ recs, unsigned long long* __restrict__ morton,
                                  unsigned int* __restrict__ cursor, unsigned long long cap)
{
  const unsigned long long n_slots = m.mask + 1;
  for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_slots;
       i += (unsigned long long)gridDim.x * blockDim.x)
  {
    const Slot s = m.slots[i];
    if (s.key == 0ull)
      continue;
    const unsigned int dst = atomicAdd(cursor, 1u);
    if (dst < cap)
    {
      const unsigned long long k = s.key & ~kValidBit;
      recs[dst] = Record{ k, s.counts };
      if (morton)
        morton[dst] = morton48(k);
    }
  }
}

__global__ void map_count_occupied_kernel(MapView m, unsigned int* __restrict__ n_occ)
{
  const unsigned long long n_slots = m.mask + 1;
  unsigned int mine = 0;
  for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_slots;
       i += (unsigned long long)gridDim.x * blockDim.x)
  {
    const Slot s = m.slots[i];
    mine += (s.key != 0ull && (s.counts >> 32) != 0ull) ? 1u : 0u;
  }
  for (int o = 16; o > 0; o >>= 1)
    mine += __shfl_xor_sync(0xffffffffu, mine, o);
  if ((threadIdx.x & 31) == 0 && mine)
    atomicAdd(n_occ, mine);
}

// [0] occupied voxels, [1] sum of count0, [2] sum of count1 over the live voxels
__global__ void map_stats_kernel(MapView m, unsigned long long* __restrict__ acc)
{
  const unsigned int n = *m.n_used;
  unsigned long long occ = 0, s0 = 0, s1 = 0;
  for (unsigned int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
  {
    const unsigned long long c = m.slots[m.live[i]].counts;
    s0 += c & 0xffffffffull;
    s1 += c >> 32;
    occ += (c >> 32) ? 1ull : 0ull;
  }
  for (int o = 16; o > 0; o >>= 1)
  {
    occ += __shfl_xor_sync(0xffffffffu, occ, o);
    s0 += __shfl_xor_sync(0xffffffffu, s0, o);
    s1 += __shfl_xor_sync(0xffffffffu, s1, o);
  }
  if ((threadIdx.x & 31) == 0 && (s0 | s1))
  {
    atomicAdd(&acc[0], occ);
    atomicAdd(&acc[1], s0);
    atomicAdd(&acc[2], s1);
  }
}

// after the sort: idx[i] = position of the i-th record in Morton order
__global__ void map_emit_kernel(const Record* __restrict__ recs, const unsigned int* __restrict__ idx, unsigned int n,
                                unsigned long long* __restrict__ key48, unsigned int* __restrict__ c0,
                                unsigned int* __restrict__ c1, float* __restrict__ logodds, float hit, float clamp_max)
{
  const unsigned int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n)
    return;
  const Record r = recs[idx[i]];
  const unsigned int n0 = (unsigned int)(r.counts & 0xffffffffull), n1 = (unsigned int)(r.counts >> 32);
  if (key48)
    key48[i] = r.key48;
  if (c0)
    c0[i] = n0;
  if (c1)
    c1[i] = n1;
  if (logodds)
  {
    // replay octomap's float updates: v += hit, clamp, stop once clamped (updateNode early return)
    float v = 0.0f;
    const unsigned int steps = n1 < 8u ? n1 : 8u;
    for (unsigned int s = 0; s < steps; s++)
    {
      if (v >= clamp_max)
        break;
      v = __fadd_rn(v, hit);
      if (v > clamp_max)
        v = clamp_max;
    }
    logodds[i] = v;
  }
}

__global__ void iota_kernel(unsigned int* __restrict__ v, unsigned int n)
{
  const unsigned int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n)
    v[i] = i;
}

__global__ void map_merge_records_kernel(MapView m, const Record* __restrict__ recs, unsigned long long n)
{
  for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (unsigned long long)gridDim.x * blockDim.x)
  {
    const Record r = recs[i];
    map_add(m, r.key48 & 0xffffffffffffull, r.counts);
  }
}

__global__ void map_merge_soa_kernel(MapView m, const unsigned long long* __restrict__ key48,
                                     const unsigned int* __restrict__ c0, const unsigned int* __restrict__ c1,
                                     unsigned long long n)
{
  for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (unsigned long long)gridDim.x * blockDim.x)
    map_add(m, key48[i] & 0xffffffffffffull,
            (unsigned long long)(c0 ? c0[i] : 0u) | ((unsigned long long)(c1 ? c1[i] : 0u) << 32));
}

// owner = mix64(key48) % n_owners; two passes over the table: histogram, then scatter into owner-grouped records
__global__ void owner_hist_kernel(MapView m, unsigned int n_owners, unsigned int* __restrict__ hist)
{
  const unsigned long long n_slots = m.mask + 1;
  for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_slots;
       i += (unsigned long long)gridDim.x * blockDim.x)
  {
    const unsigned long long k = m.slots[i].key;
    if (k)
      atomicAdd(&hist[mix64(k & ~kValidBit) % n_owners], 1u);
  }
}

__global__ void owner_scatter_kernel(MapView m, unsigned int n_owners, const unsigned int* __restrict__ start,
                                     unsigned int* __restrict__ cursor, Record* __restrict__ recs, unsigned long long cap)
{
  const unsigned long long n_slots = m.mask + 1;
  for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_slots;
       i += (unsigned long long)gridDim.x * blockDim.x)
  {
    const Slot s = m.slots[i];
    if (!s.key)
      continue;
    const unsigned long long k = s.key & ~kValidBit;
    const unsigned int o = (unsigned int)(mix64(k) % n_owners);
    const unsigned long long dst = (unsigned long long)start[o] + atomicAdd(&cursor[o], 1u);
    if (dst < cap)
      recs[dst] = Record{ k, s.counts };
  }
}

constexpr int kCounterWords = 128;
constexpr int kEpochFlagWord = 64;   // [64, 80): exchange epoch flags, one per peer rank (never reset)
constexpr int kClearDoneWord = 127;  // CTAs done in map_clear_kernel / map_push_to_owners_kernel (back to 0 when they end)

// ---- peer-memory exchange: every used slot of the local table is added straight into its owner's table ----------------
// The owner tables may live on other GPUs (CUDA IPC mappings reached over NVLink), so the insert uses system-scope
// atomics: a 64-bit CAS claims the slot at the owner's L2, a 64-bit add merges the two counters.
struct OwnerViews
{
  MapView v[16];
};

// adds `add` to voxel key48 of a (possibly remote) table; returns true if this call claimed a new slot (*slot_out),
// in which case the caller appends it to the table's live list
__device__ __forceinline__ bool map_add_system(const MapView& m, unsigned long long key48, unsigned long long add,
                                               unsigned int* slot_out)
{
  const unsigned long long want = key48 | kValidBit;
  unsigned long long slot = slot_hash(key48) & m.mask;
#pragma unroll 1
  for (unsigned int probe = 0; probe < 65536u; ++probe)
  {
    Slot* s = m.slots + slot;
    const unsigned long long k = atomicCAS_system(&s->key, 0ull, want);  // one round trip: claims or reads the key
    if (k == 0ull || k == want)
    {
      atomicAdd_system(&s->counts, add);
      *slot_out = (unsigned int)slot;
      return k == 0ull;
    }
    slot = (slot + 1) & m.mask;
  }
  *m.overflow = 1u;
  return false;
}

// Walks the live voxels of the local table (not its slots: a table sized for the worst case is mostly empty) and adds
// each to its owner's table.  New claims are appended to the owner's live list with ONE system-scope atomic per warp and
// owner (the lanes that claimed a slot in the same table are counted with a ballot): a returning atomic per voxel on one
// remote address was the dominant cost of the exchange.  With clear_local the kernel also leaves the local table empty
// (each slot is zeroed once it has been read, the last CTA resets the counters), which saves the separate clear.
__global__ void __launch_bounds__(256) map_push_to_owners_kernel(MapView local, OwnerViews owners, unsigned int n_owners,
                                                                int clear_local, unsigned int* __restrict__ local_counters)
{
  const unsigned int n = *local.n_used;
  const unsigned int lane = threadIdx.x & 31u;
  const unsigned int n_round = (n + 31u) & ~31u;  // whole warps stay converged for the ballots
  for (unsigned int i = blockIdx.x * blockDim.x + threadIdx.x; i < n_round; i += gridDim.x * blockDim.x)
  {
    bool claimed = false;
    unsigned int owner = 0, new_slot = 0;
    if (i < n)
    {
      const unsigned int ls = local.live[i];
      const Slot s = local.slots[ls];
      if (clear_local)
        reinterpret_cast<uint4*>(local.slots)[ls] = make_uint4(0u, 0u, 0u, 0u);
      const unsigned long long k = s.key & ~kValidBit;
      owner = (unsigned int)(mix64(k) % n_owners);
      claimed = map_add_system(owners.v[owner], k, s.counts, &new_slot);
    }
    unsigned int pending = __ballot_sync(0xffffffffu, claimed);
    while (pending)
    {
      const int leader = __ffs(pending) - 1;
      const unsigned int o = __shfl_sync(0xffffffffu, owner, leader);
      const unsigned int grp = __ballot_sync(0xffffffffu, claimed && owner == o);
      unsigned int base = 0;
      if ((int)lane == leader)
        base = atomicAdd_system(owners.v[o].n_used, (unsigned int)__popc(grp));
      base = __shfl_sync(0xffffffffu, base, leader);
      if (claimed && owner == o)
        owners.v[o].live[base + __popc(grp & ((1u << lane) - 1u))] = new_slot;
      pending &= ~grp;
    }
  }
  if (clear_local)
  {
    __syncthreads();
    if (threadIdx.x == 0)
    {
      __threadfence();
      if (atomicAdd(&local_counters[kClearDoneWord], 1u) == gridDim.x - 1)
      {
        for (int c = 0; c < 64; c++)
          local_counters[c] = 0u;
        local_counters[kClearDoneWord] = 0u;
      }
    }
  }
}

// ---- mailbox form of the push: no atomics cross NVLink -----------------------------------------------------------------------
// Voxels owned by this rank are added to its own table with ordinary device-scope atomics.  Every other voxel becomes a
// 16-byte record {key48, count0 | count1 << 32} stored (plain, posted stores) into this rank's segment of the owner's inbox;
// the slot in the segment comes from a LOCAL cursor, advanced once per warp and owner.  When the last CTA is done it
// publishes the record counts in the owners' inbox headers.  After the barrier each owner merges its inbox with
// map_merge_inbox_kernel, again with local atomics only.  (System-scope CAS / RED on remote tables cost ~1 us per thousand
// voxels and collapse when several ranks update the same voxels -- overlapping scans: 1.4 ms per exchange at 8 GPUs.)
constexpr int kCursorWord = 80;  // [80, 96): per-owner record cursors of map_push_mailbox_kernel (back to 0 when it ends)

struct MailboxPtrs
{
  Record* seg[16];               // this rank's segment in owner o's inbox (peer memory); unused for o == rank
  unsigned long long* count[16]; // this rank's count word in owner o's inbox header
};

__global__ void __launch_bounds__(256) map_push_mailbox_kernel(MapView local, MapView own, unsigned int rank, unsigned int n_owners,
                                                              MailboxPtrs mb, unsigned long long seg_cap, int clear_local,
                                                              unsigned int* __restrict__ local_counters)
{
  const unsigned int n = *local.n_used;
  const unsigned int lane = threadIdx.x & 31u;
  const unsigned int n_round = (n + 31u) & ~31u;  // whole warps stay converged for the ballots
  unsigned int* cursor = local_counters + kCursorWord;
  for (unsigned int i = blockIdx.x * blockDim.x + threadIdx.x; i < n_round; i += gridDim.x * blockDim.x)
  {
    bool remote = false, claimed = false;
    unsigned int owner = 0, own_slot = 0;
    Record rec{ 0ull, 0ull };
    if (i < n)
    {
      const unsigned int ls = local.live[i];
      const Slot s = local.slots[ls];
      if (clear_local)
        reinterpret_cast<uint4*>(local.slots)[ls] = make_uint4(0u, 0u, 0u, 0u);
      rec.key48 = s.key & ~kValidBit;
      rec.counts = s.counts;
      owner = (unsigned int)(mix64(rec.key48) % n_owners);
      if (owner == rank)
        claimed = map_add_claim(own, rec.key48, rec.counts, &own_slot);
      else
        remote = true;
    }
    live_append_warp(own, claimed, own_slot);
    // one segment slot per remote record: the lanes of a warp that send to the same owner are grouped with match.any, each
    // group's first lane advances that owner's cursor once, all groups at the same time (one atomic round trip per warp)
    const unsigned int peers = __match_any_sync(0xffffffffu, remote ? owner : 0xffffffffu);
    const int leader = __ffs(peers) - 1;
    unsigned int base = 0;
    if (remote && (int)lane == leader)
      base = atomicAdd(&cursor[owner], (unsigned int)__popc(peers));
    base = __shfl_sync(0xffffffffu, base, leader);
    if (remote)
    {
      const unsigned long long pos = (unsigned long long)base + __popc(peers & ((1u << lane) - 1u));
      if (pos < seg_cap)
        *reinterpret_cast<ulonglong2*>(mb.seg[owner] + pos) = make_ulonglong2(rec.key48, rec.counts);
      else
        *local.overflow = 1u;  // more voxels for one owner than the segment holds: reported by pcs_map_size
    }
  }
  // Visibility of the records at the owners is the barrier's business: it follows this kernel in the stream and publishes
  // the epoch with a system-scope release, which is cumulative over everything that happened before it.  (A system-scope
  // fence per CTA here cost ~15 us.)  Only the cursors need ordering inside this kernel: device scope.
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0)
  {
    if (atomicAdd(&local_counters[kClearDoneWord], 1u) == gridDim.x - 1)
    {
      __threadfence();
      for (unsigned int o = 0; o < n_owners; o++)
      {
        const unsigned int c = *reinterpret_cast<volatile unsigned int*>(&cursor[o]);
        if (o != rank)
          *reinterpret_cast<volatile unsigned long long*>(mb.count[o]) = c < seg_cap ? c : seg_cap;
        cursor[o] = 0u;
      }
      if (clear_local)
        for (int c = 0; c < 64; c++)
          local_counters[c] = 0u;
      local_counters[kClearDoneWord] = 0u;
    }
  }
}

// blockIdx.y = source rank; the records its push left in this rank's inbox are added to the owned table
__global__ void __launch_bounds__(256) map_merge_inbox_kernel(MapView own, const unsigned long long* __restrict__ header,
                                                             const Record* __restrict__ segs, unsigned long long seg_cap,
                                                             unsigned int rank)
{
  const unsigned int src = blockIdx.y;
  if (src == rank)
    return;
  const unsigned long long n = __ldcg(&header[src]);
  const Record* seg = segs + (unsigned long long)src * seg_cap;
  const unsigned long long n_round = (n + 31ull) & ~31ull;  // whole warps stay converged for the ballot
  for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_round;
       i += (unsigned long long)gridDim.x * blockDim.x)
  {
    bool claimed = false;
    unsigned int slot = 0;
    if (i < n)
    {
      const ulonglong2 r = __ldcg(reinterpret_cast<const ulonglong2*>(seg + i));  // written by a peer: read at L2
      claimed = map_add_claim(own, r.x, r.y, &slot);
    }
    live_append_warp(own, claimed, slot);
  }
}

// Clears a table in time proportional to its live voxels: the claimed slots are zeroed through the live list (or the
// whole table linearly when more than an eighth of it is in use), then the last CTA to finish resets the counters.
__global__ void __launch_bounds__(256) map_clear_kernel(MapView m, unsigned int* __restrict__ counters)
{
  const unsigned int n = *m.n_used;
  const unsigned long long n_slots = m.mask + 1;
  const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
  const unsigned long long first = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
  if ((unsigned long long)n > n_slots / 8 || *m.overflow)
  {
    uint4* p = reinterpret_cast<uint4*>(m.slots);
    for (unsigned long long i = first; i < n_slots; i += stride)
      p[i] = make_uint4(0u, 0u, 0u, 0u);
  }
  else
    for (unsigned long long i = first; i < n; i += stride)
      reinterpret_cast<uint4*>(m.slots)[m.live[i]] = make_uint4(0u, 0u, 0u, 0u);
  __syncthreads();
  if (threadIdx.x == 0)
  {
    __threadfence();
    if (atomicAdd(&counters[kClearDoneWord], 1u) == gridDim.x - 1)
    {
      for (int i = 0; i < 64; i++)
        counters[i] = 0u;
      counters[kClearDoneWord] = 0u;
    }
  }
}

static unsigned int table_grid(pcs_ctx* ctx, unsigned long long n)
{
  unsigned long long b = (n + 255) / 256;
  const unsigned long long mx = (unsigned long long)ctx->sm_count * 8ull;
  return (unsigned int)(b < mx ? (b ? b : 1) : mx);
}

// one allocation per map: slots (16 B each) | live list (4 B per slot) | counters (512 B); peers map it with one IPC handle
static size_t map_block_bytes(uint64_t n_slots)
{
  return (size_t)n_slots * sizeof(Slot) + (size_t)n_slots * sizeof(unsigned int) + kCounterWords * sizeof(unsigned int);
}

static void map_block_views(pcs_map* m, double resolution)
{
  char* base = (char*)m->block;
  m->view.slots = (Slot*)base;
  m->view.live = (unsigned int*)(base + (size_t)m->n_slots * sizeof(Slot));
  m->d_counters = (unsigned int*)(base + (size_t)m->n_slots * (sizeof(Slot) + sizeof(unsigned int)));
  m->view.mask = m->n_slots - 1;
  m->view.n_used = m->d_counters;
  m->view.overflow = m->d_counters + 1;
  m->view.inv_res = 1.0 / resolution;  // octomap: resolution_factor = 1. / resolution
}

static int read_counters(pcs_ctx* ctx, pcs_map* map, unsigned int out[2])
{
  void* h = nullptr;
  PCS_TRY(ctx_pinned(ctx, 64, &h));
  PCS_CUDA(cudaMemcpyAsync(h, map->d_counters, 2 * sizeof(unsigned int), cudaMemcpyDeviceToHost, ctx->stream));
  PCS_CUDA(cudaStreamSynchronize(ctx->stream));
  out[0] = ((unsigned int*)h)[0];
  out[1] = ((unsigned int*)h)[1];
  if (out[1] || out[0] > map->capacity_voxels)
  {
    set_error("voxel map overflow: %u distinct voxels, capacity %llu", out[0], (unsigned long long)map->capacity_voxels);
    return PCS_ERR_CAPACITY;
  }
  return PCS_OK;
}
}  // namespace pcs

using namespace pcs;

extern "C" {

uint64_t pcs_mix64(uint64_t x)
{
  x ^= x >> 33;
  x *= 0xff51afd7ed558ccdULL;
  x ^= x >> 33;
  x *= 0xc4ceb9fe1a85ec53ULL;
  x ^= x >> 33;
  return x;
}

int pcs_map_create(pcs_ctx* ctx, double resolution, uint64_t capacity_voxels, pcs_map** out)
{
  PCS_TRY(ctx_bind(ctx));
  PCS_CHECK_ARG(out != nullptr, "map out pointer is NULL");
  *out = nullptr;
  PCS_CHECK_ARG(resolution > 0.0 && isfinite(resolution), "resolution must be positive");
  PCS_CHECK_ARG(capacity_voxels > 0 && capacity_voxels <= (1ull << 31), "capacity_voxels out of range");
  uint64_t n_slots = 1024;
  while (n_slots < 2 * capacity_voxels)  // load factor <= 0.5
    n_slots <<= 1;
  pcs_map* m = new (std::nothrow) pcs_map();
  PCS_CHECK_ARG(m != nullptr, "out of host memory");
  m->resolution = resolution;
  m->capacity_voxels = capacity_voxels;
  m->n_slots = n_slots;
  m->is_peer = false;
  m->d_counters = nullptr;
  m->view.slots = nullptr;
  m->block = nullptr;
  const size_t bytes = map_block_bytes(n_slots);
  cudaError_t e = cudaMalloc(&m->block, bytes);
  if (e != cudaSuccess)
  {
    delete m;
    return cuda_fail(e, "cudaMalloc(voxel map)", __FILE__, __LINE__);
  }
  map_block_views(m, resolution);
  e = cudaMemsetAsync(m->block, 0, bytes, ctx->stream);  // the only full memset: clears walk the live list
  if (e != cudaSuccess)
  {
    cudaFree(m->block);
    delete m;
    return cuda_fail(e, "cudaMemsetAsync(voxel map)", __FILE__, __LINE__);
  }
  *out = m;
  return PCS_OK;
}

int pcs_map_destroy(pcs_ctx* ctx, pcs_map* map)
{
  if (!map)
    return PCS_OK;
  PCS_TRY(ctx_bind(ctx));
  PCS_CUDA(cudaStreamSynchronize(ctx->stream));
  if (map->is_peer)
    cudaIpcCloseMemHandle(map->block);
  else
    cudaFree(map->block);
  delete map;
  return PCS_OK;
}

int pcs_map_clear(pcs_ctx* ctx, pcs_map* map)
{
  PCS_TRY(ctx_bind(ctx));
  PCS_CHECK_ARG(map != nullptr, "map is NULL");
  map_clear_kernel<<<(unsigned int)ctx->sm_count * 2u, 256, 0, ctx->stream>>>(map->view, map->d_counters);
  count_launch(ctx);
  return check_launch("map_clear_kernel");
}

int pcs_map_size(pcs_ctx* ctx, pcs_map* map, uint64_t* n_voxels, uint64_t* n_occupied)
{
  PCS_TRY(ctx_bind(ctx));
  PCS_CHECK_ARG(map != nullptr, "map is NULL");
  if (n_occupied)
  {
    PCS_CUDA(cudaMemsetAsync(map->d_counters + 2, 0, sizeof(unsigned int), ctx->stream));
    map_count_occupied_kernel<<<table_grid(ctx, map->n_slots), 256, 0, ctx->stream>>>(map->view, map->d_counters + 2);
    count_launch(ctx);
    PCS_TRY(check_launch("map_count_occupied_kernel"));
    void* h = nullptr;
    PCS_TRY(ctx_pinned(ctx, 64, &h));
    PCS_CUDA(cudaMemcpyAsync((char*)h + 16, map->d_counters + 2, 4, cudaMemcpyDeviceToHost, ctx->stream));
    PCS_CUDA(cudaStreamSynchronize(ctx->stream));
    *n_occupied = *(unsigned int*)((char*)h + 16);
  }
  unsigned int c[2];
  PCS_TRY(read_counters(ctx, map, c));
  if (n_voxels)
    *n_voxels = c[0];
  return PCS_OK;
}

int pcs_map_get_stats(pcs_ctx* ctx, pcs_map* map, pcs_map_stats* out)
{
  PCS_TRY(ctx_bind(ctx));
  PCS_CHECK_ARG(map && out, "NULL argument");
  memset(out, 0, sizeof(*out));
  void* scr = nullptr;
  PCS_TRY(ctx_scratch(ctx, 64, &scr));
  unsigned long long* acc = (unsigned long long*)scr;
  PCS_CUDA(cudaMemsetAsync(acc, 0, 32, ctx->stream));
  map_stats_kernel<<<(unsigned int)ctx->sm_count * 2u, 256, 0, ctx->stream>>>(map->view, acc);
  count_launch(ctx);
  PCS_TRY(check_launch("map_stats_kernel"));
  void* h = nullptr;
  PCS_TRY(ctx_pinned(ctx, 128, &h));
  PCS_CUDA(cudaMemcpyAsync((char*)h + 64, acc, 32, cudaMemcpyDeviceToHost, ctx->stream));  // read_counters uses the first bytes
  unsigned int c[2];
  PCS_TRY(read_counters(ctx, map, c));  // synchronises the stream
  const unsigned long long* v = (const unsigned long long*)((char*)h + 64);
  out->voxels = c[0];
  out->occupied = v[0];
  out->points_rejected = v[1];
  out->points_labelled = v[2];
  out->table_slots = map->n_slots;
  out->load_factor = (double)c[0] / (double)map->n_slots;
  return PCS_OK;
}

int pcs_map_export(pcs_ctx* ctx, pcs_map* map, uint64_t cap, uint64_t* key48, uint32_t* count0, uint32_t* count1,
                   float* logodds, uint64_t* n_out)
{
  PCS_TRY(ctx_bind(ctx));
  PCS_CHECK_ARG(map && n_out, "NULL argument");
  unsigned int c[2];
  PCS_TRY(read_counters(ctx, map, c));
  const unsigned int n = c[0];
  *n_out = n;
  if (n == 0)
    return PCS_OK;
  if (cap < n)
  {
    if (key48 || count0 || count1 || logodds)
    {
      set_error("export buffers hold %llu records, map has %u", (unsigned long long)cap, n);
      return PCS_ERR_CAPACITY;
    }
    return PCS_OK;  // pure size query
  }
  // scratch: recs | morton_in | morton_out | idx_in | idx_out | cub temp | outputs
  size_t cub_bytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, cub_bytes, (unsigned long long*)nullptr, (unsigned long long*)nullptr,
                                  (unsigned int*)nullptr, (unsigned int*)nullptr, (int)n, 0, 48, ctx->stream);
  auto al = [](size_t b) { return (b + 255) & ~(size_t)255; };
  const size_t o_recs = 0, o_min = o_recs + al((size_t)n * sizeof(Record)), o_mout = o_min + al((size_t)n * 8),
               o_iin = o_mout + al((size_t)n * 8), o_iout = o_iin + al((size_t)n * 4), o_cub = o_iout + al((size_t)n * 4),
               o_key = o_cub + al(cub_bytes), o_c0 = o_key + al((size_t)n * 8), o_c1 = o_c0 + al((size_t)n * 4),
               o_lo = o_c1 + al((size_t)n * 4), total = o_lo + al((size_t)n * 4);
  void* scr = nullptr;
  PCS_TRY(ctx_scratch(ctx, total, &scr));
  char* base = (char*)scr;
  Record* d_recs = (Record*)(base + o_recs);
  unsigned long long *d_min = (unsigned long long*)(base + o_min), *d_mout = (unsigned long long*)(base + o_mout);
  unsigned int *d_iin = (unsigned int*)(base + o_iin), *d_iout = (unsigned int*)(base + o_iout);
  PCS_CUDA(cudaMemsetAsync(map->d_counters + 3, 0, sizeof(unsigned int), ctx->stream));
  map_gather_kernel<<<table_grid(ctx, map->n_slots), 256, 0, ctx->stream>>>(map->view, d_recs, d_min, map->d_counters + 3, n);
  iota_kernel<<<(n + 255) / 256, 256, 0, ctx->stream>>>(d_iin, n);
  count_launch(ctx, 2);
  PCS_TRY(check_launch("map_gather_kernel"));
  PCS_CUDA(cub::DeviceRadixSort::SortPairs(base + o_cub, cub_bytes, d_min, d_mout, d_iin, d_iout, (int)n, 0, 48, ctx->stream));
  const float hit = (float)log(0.7 / (1.0 - 0.7)), cmax = (float)log(0.971 / (1.0 - 0.971));
  map_emit_kernel<<<(n + 255) / 256, 256, 0, ctx->stream>>>(
      d_recs, d_iout, n, key48 ? (unsigned long long*)(base + o_key) : nullptr, count0 ? (unsigned int*)(base + o_c0) : nullptr,
      count1 ? (unsigned int*)(base + o_c1) : nullptr, logodds ? (float*)(base + o_lo) : nullptr, hit, cmax);
  count_launch(ctx);
  PCS_TRY(check_launch("map_emit_kernel"));
  if (key48)
    PCS_CUDA(cudaMemcpyAsync(key48, base + o_key, (size_t)n * 8, cudaMemcpyDeviceToHost, ctx->stream));
  if (count0)
    PCS_CUDA(cudaMemcpyAsync(count0, base + o_c0, (size_t)n * 4, cudaMemcpyDeviceToHost, ctx->stream));
  if (count1)
    PCS_CUDA(cudaMemcpyAsync(count1, base + o_c1, (size_t)n * 4, cudaMemcpyDeviceToHost, ctx->stream));
  if (logodds)
    PCS_CUDA(cudaMemcpyAsync(logodds, base + o_lo, (size_t)n * 4, cudaMemcpyDeviceToHost, ctx->stream));
  PCS_CUDA(cudaStreamSynchronize(ctx->stream));
  ctx->stats.bytes_d2h += (size_t)n * 20;
  return PCS_OK;
}

int pcs_map_merge_records_dev(pcs_ctx* ctx, pcs_map* map, const void* records_dev, uint64_t n)
{
  PCS_TRY(ctx_bind(ctx));
  PCS_CHECK_ARG(map != nullptr, "map is NULL");
  if (n == 0)
    return PCS_OK;
  PCS_CHECK_ARG(records_dev != nullptr, "records is NULL");
  map_merge_records_kernel<<<table_grid(ctx, n), 256, 0, ctx->stream>>>(map->view, (const Record*)records_dev, n);
  count_launch(ctx);
  return check_launch("map_merge_records_kernel");
}

int pcs_map_merge(pcs_ctx* ctx, pcs_map* map, const uint64_t* key48, const uint32_t* count0, const uint32_t* count1,
                  uint64_t n)
{
  PCS_TRY(ctx_bind(ctx));
  PCS_CHECK_ARG(map != nullptr, "map is NULL");
  if (n == 0)
    return PCS_OK;
  PCS_CHECK_ARG(key48 != nullptr, "key48 is NULL");
  void* d = nullptr;
  PCS_TRY(ctx_staging(ctx, 0, n * 16 + 512, &d));
  unsigned long long* d_k = (unsigned long long*)d;
  unsigned int* d_c0 = (unsigned int*)((char*)d + n * 8);
  unsigned int* d_c1 = (unsigned int*)((char*)d + n * 12);
  PCS_CUDA(cudaMemcpyAsync(d_k, key48, n * 8, cudaMemcpyHostToDevice, ctx->stream));
  if (count0)
    PCS_CUDA(cudaMemcpyAsync(d_c0, count0, n * 4, cudaMemcpyHostToDevice, ctx->stream));
  if (count1)
    PCS_CUDA(cudaMemcpyAsync(d_c1, count1, n * 4, cudaMemcpyHostToDevice, ctx->stream));
  map_merge_soa_kernel<<<table_grid(ctx, n), 256, 0, ctx->stream>>>(map->view, d_k, count0 ? d_c0 : nullptr,
                                                                  count1 ? d_c1 : nullptr, n);
  count_launch(ctx);
  PCS_TRY(check_launch("map_merge_soa_kernel"));
  PCS_CUDA(cudaStreamSynchronize(ctx->stream));
  ctx->stats.bytes_h2d += n * 16;
  return PCS_OK;
}

int pcs_map_ipc_export(pcs_ctx* ctx, pcs_map* map, void* handle128)
{
  PCS_TRY(ctx_bind(ctx));
  PCS_CHECK_ARG(map && handle128, "NULL argument");
  PCS_CHECK_ARG(!map->is_peer, "cannot re-export a peer map");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  cudaIpcMemHandle_t h[2];
  memset(h, 0, sizeof(h));  // the second half of the 128 bytes is reserved (the map is one allocation)
  PCS_CUDA(cudaIpcGetMemHandle(&h[0], map->block));
  memcpy(handle128, h, sizeof(h));
  return PCS_OK;
}

int pcs_map_ipc_open(pcs_ctx* ctx, const void* handle128, double resolution, uint64_t capacity_voxels, pcs_map** out)
{
  PCS_TRY(ctx_bind(ctx));
  PCS_CHECK_ARG(handle128 && out, "NULL argument");
  *out = nullptr;
  PCS_CHECK_ARG(resolution > 0.0 && capacity_voxels > 0 && capacity_voxels <= (1ull << 31), "bad map geometry");
  uint64_t n_slots = 1024;
  while (n_slots < 2 * capacity_voxels)  // must mirror pcs_map_create
    n_slots <<= 1;
  cudaIpcMemHandle_t h[2];
  memcpy(h, handle128, sizeof(h));
  void* block = nullptr;
  PCS_CUDA(cudaIpcOpenMemHandle(&block, h[0], cudaIpcMemLazyEnablePeerAccess));
  pcs_map* m = new (std::nothrow) pcs_map();
  if (!m)
  {
    cudaIpcCloseMemHandle(block);
    set_error("out of host memory");
    return PCS_ERR_BAD_ARG;
  }
  m->resolution = resolution;
  m->capacity_voxels = capacity_voxels;
  m->n_slots = n_slots;
  m->is_peer = true;
  m->block = block;
  map_block_views(m, resolution);
  *out = m;
  return PCS_OK;
}

// ---- stream-ordered flags: the barrier of the peer-memory exchange without a collective and without a spinning kernel ----
typedef CUresult (*stream_write32_fn)(CUstream, CUdeviceptr, cuuint32_t, unsigned int);
typedef CUresult (*stream_wait32_fn)(CUstream, CUdeviceptr, cuuint32_t, unsigned int);

static int stream_memops(stream_write32_fn* wr, stream_wait32_fn* wt)
{
  static stream_write32_fn s_wr = nullptr;
  static stream_wait32_fn s_wt = nullptr;
  if (!s_wr || !s_wt)
  {
    cudaDriverEntryPointQueryResult q1, q2;
    void *p1 = nullptr, *p2 = nullptr;
    PCS_CUDA(cudaGetDriverEntryPoint("cuStreamWriteValue32", &p1, cudaEnableDefault, &q1));
    PCS_CUDA(cudaGetDriverEntryPoint("cuStreamWaitValue32", &p2, cudaEnableDefault, &q2));
    if (q1 != cudaDriverEntryPointSuccess || q2 != cudaDriverEntryPointSuccess || !p1 || !p2)
    {
      set_error("the driver does not offer cuStreamWriteValue32 / cuStreamWaitValue32");
      return PCS_ERR_CUDA;
    }
    s_wr = (stream_write32_fn)p1;
    s_wt = (stream_wait32_fn)p2;
  }
  *wr = s_wr;
  *wt = s_wt;
  return PCS_OK;
}

int pcs_map_epoch_flag_dev(pcs_map* map, int index, void** flag_dev)
{
  PCS_CHECK_ARG(map && flag_dev, "NULL argument");
  PCS_CHECK_ARG(index >= 0 && index < 16, "flag index must be in [0, 16)");
  *flag_dev = map->d_counters + kEpochFlagWord + index;
  return PCS_OK;
}

int pcs_stream_write_u32_dev(pcs_ctx* ctx, void* flag_dev, uint32_t value)
{
  PCS_TRY(ctx_bind(ctx));
  PCS_CHECK_ARG(flag_dev != nullptr && ((uintptr_t)flag_dev & 3u) == 0, "flag pointer must be a 4-byte aligned device address");
  stream_write32_fn wr;
  stream_wait32_fn wt;
  PCS_TRY(stream_memops(&wr, &wt));
  const CUresult r = wr((CUstream)ctx->stream, (CUdeviceptr)(uintptr_t)flag_dev, value, CU_STREAM_WRITE_VALUE_DEFAULT);
  if (r != CUDA_SUCCESS)
  {
    set_error("cuStreamWriteValue32 failed with CUresult %d", (int)r);
    return PCS_ERR_CUDA;
  }
  return PCS_OK;
}

int pcs_stream_wait_geq_u32_dev(pcs_ctx* ctx, void* flag_dev, uint32_t value)
{
  PCS_TRY(ctx_bind(ctx));
  PCS_CHECK_ARG(flag_dev != nullptr && ((uintptr_t)flag_dev & 3u) == 0, "flag pointer must be a 4-byte aligned device address");
  stream_write32_fn wr;
  stream_wait32_fn wt;
  PCS_TRY(stream_memops(&wr, &wt));
  const CUresult r = wt((CUstream)ctx->stream, (CUdeviceptr)(uintptr_t)flag_dev, value, CU_STREAM_WAIT_VALUE_GEQ);
  if (r != CUDA_SUCCESS)
  {
    set_error("cuStreamWaitValue32 failed with CUresult %d", (int)r);
    return PCS_ERR_CUDA;
  }
  return PCS_OK;
}

int pcs_ipc_export_dev(pcs_ctx* ctx, void* dptr, void* handle64)
{
  PCS_TRY(ctx_bind(ctx));
  PCS_CHECK_ARG(dptr && handle64, "NULL argument");
  cudaIpcMemHandle_t h;
  PCS_CUDA(cudaIpcGetMemHandle(&h, dptr));
  memcpy(handle64, &h, sizeof(h));
  return PCS_OK;
}

int pcs_ipc_open_dev(pcs_ctx* ctx, const void* handle64, void** dptr)
{
  PCS_TRY(ctx_bind(ctx));
  PCS_CHECK_ARG(dptr && handle64, "NULL argument");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, sizeof(h));
  PCS_CUDA(cudaIpcOpenMemHandle(dptr, h, cudaIpcMemLazyEnablePeerAccess));
  return PCS_OK;
}

int pcs_ipc_close_dev(pcs_ctx* ctx, void* dptr)
{
  PCS_TRY(ctx_bind(ctx));
  if (dptr)
    PCS_CUDA(cudaIpcCloseMemHandle(dptr));
  return PCS_OK;
}

uint64_t pcs_map_inbox_bytes(uint64_t capacity_voxels, int n_owners)
{
  return 256ull + (uint64_t)n_owners * capacity_voxels * sizeof(Record);
}

int pcs_map_push_mailbox_dev(pcs_ctx* ctx, pcs_map* local, pcs_map* own, int rank, int n_owners, void* const* peer_inboxes,
                             uint64_t seg_capacity, int clear_local)
{
  PCS_TRY(ctx_bind(ctx));
  PCS_CHECK_ARG(local && own && peer_inboxes, "NULL argument");
  PCS_CHECK_ARG(n_owners >= 1 && n_owners <= 16 && rank >= 0 && rank < n_owners, "rank / n_owners out of range");
  PCS_CHECK_ARG(own != local && own->resolution == local->resolution, "owned map must be a different map of the same resolution");
  PCS_CHECK_ARG(seg_capacity > 0, "seg_capacity must be positive");
  MailboxPtrs mb;
  for (int o = 0; o < 16; o++)
  {
    mb.seg[o] = nullptr;
    mb.count[o] = nullptr;
    if (o < n_owners && o != rank)
    {
      PCS_CHECK_ARG(peer_inboxes[o] != nullptr, "peer inbox is NULL");
      // owner o's inbox: header word [rank], then segment [rank]
      mb.count[o] = (unsigned long long*)peer_inboxes[o] + rank;
      mb.seg[o] = (Record*)((char*)peer_inboxes[o] + 256) + (unsigned long long)rank * seg_capacity;
    }
  }
  unsigned int grid = table_grid(ctx, local->capacity_voxels);
  if (grid > (unsigned int)ctx->sm_count * 2u)
    grid = (unsigned int)ctx->sm_count * 2u;
  map_push_mailbox_kernel<<<grid, 256, 0, ctx->stream>>>(local->view, own->view, (unsigned int)rank, (unsigned int)n_owners, mb,
                                                         seg_capacity, clear_local ? 1 : 0, local->d_counters);
  count_launch(ctx);
  return check_launch("map_push_mailbox_kernel");
}

int pcs_map_merge_inbox_dev(pcs_ctx* ctx, pcs_map* own, const void* inbox, int rank, int n_owners, uint64_t seg_capacity)
{
  PCS_TRY(ctx_bind(ctx));
  PCS_CHECK_ARG(own && inbox, "NULL argument");
  PCS_CHECK_ARG(n_owners >= 1 && n_owners <= 16 && rank >= 0 && rank < n_owners, "rank / n_owners out of range");
  if (n_owners == 1)
    return PCS_OK;
  dim3 grid((unsigned int)ctx->sm_count, (unsigned int)n_owners);
  map_merge_inbox_kernel<<<grid, 256, 0, ctx->stream>>>(own->view, (const unsigned long long*)inbox,
                                                        (const Record*)((const char*)inbox + 256), seg_capacity,
                                                        (unsigned int)rank);
  count_launch(ctx);
  return check_launch("map_merge_inbox_kernel");
}

// One small kernel as the barrier: lane i publishes this rank's epoch in peer i's word (release, system scope) and then
// waits for peer i's epoch in this rank's own word (acquire).  Every rank runs on its own GPU, so the wait cannot starve
// the kernel it waits for; a peer that never arrives (crashed process) trips the cycle budget and the kernel traps instead
// of hanging the device.
struct BarrierFlags
{
  unsigned int* theirs[16];
  unsigned int* mine[16];
};

__global__ void exchange_barrier_kernel(BarrierFlags f, int n, unsigned int epoch)
{
  const int i = (int)threadIdx.x;
  if (i >= n)
    return;
  __threadfence_system();
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(f.theirs[i]), "r"(epoch) : "memory");
  const long long t0 = clock64();
  for (;;)
  {
    unsigned int v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(f.mine[i]) : "memory");
    if ((int)(v - epoch) >= 0)
      break;
    if (clock64() - t0 > 20000000000ll)  // ~10 s: a peer is gone
      __trap();
  }
}

int pcs_exchange_barrier_dev(pcs_ctx* ctx, void* const* their_flags, void* const* my_flags, int n, uint32_t epoch)
{
  PCS_TRY(ctx_bind(ctx));
  PCS_CHECK_ARG(n >= 0 && n <= 16, "n must be in [0, 16]");
  if (n == 0)
    return PCS_OK;
  PCS_CHECK_ARG(their_flags && my_flags, "NULL argument");
  BarrierFlags f;
  for (int i = 0; i < 16; i++)
  {
    f.theirs[i] = i < n ? (unsigned int*)their_flags[i] : nullptr;
    f.mine[i] = i < n ? (unsigned int*)my_flags[i] : nullptr;
    PCS_CHECK_ARG(i >= n || (f.theirs[i] && f.mine[i]), "NULL flag pointer");
  }
  exchange_barrier_kernel<<<1, 32, 0, ctx->stream>>>(f, n, epoch);
  count_launch(ctx);
  return check_launch("exchange_barrier_kernel");
}

int pcs_map_push_to_owners_dev(pcs_ctx* ctx, pcs_map* local, pcs_map* const* owners, int n_owners, int clear_local)
{
  PCS_TRY(ctx_bind(ctx));
  PCS_CHECK_ARG(local && owners, "NULL argument");
  PCS_CHECK_ARG(n_owners >= 1 && n_owners <= 16, "n_owners must be in [1, 16]");
  OwnerViews ov;
  for (int o = 0; o < n_owners; o++)
  {
    PCS_CHECK_ARG(owners[o] != nullptr, "owner map is NULL");
    PCS_CHECK_ARG(owners[o] != local, "the local map cannot be one of the owners");
    PCS_CHECK_ARG(owners[o]->resolution == local->resolution, "owner map resolution differs");
    ov.v[o] = owners[o]->view;
  }
  // the grid is sized for the capacity (the live count stays on the device); CTAs beyond the live voxels exit at once
  unsigned int grid = table_grid(ctx, local->capacity_voxels);
  if (grid > (unsigned int)ctx->sm_count * 2u)
    grid = (unsigned int)ctx->sm_count * 2u;
  map_push_to_owners_kernel<<<grid, 256, 0, ctx->stream>>>(local->view, ov, (unsigned int)n_owners, clear_local ? 1 : 0,
                                                           local->d_counters);
  count_launch(ctx);
  return check_launch("map_push_to_owners_kernel");
}

int pcs_map_pack_by_owner_dev(pcs_ctx* ctx, pcs_map* map, int n_owners, void* records_dev, uint64_t cap_records,
                              uint64_t* owner_counts)
{
  PCS_TRY(ctx_bind(ctx));
  PCS_CHECK_ARG(map && owner_counts, "NULL argument");
  PCS_CHECK_ARG(n_owners >= 1 && n_owners <= 16, "n_owners must be in [1, 16]");
  // d_counters layout: [8..8+16) histogram, [24..24+16) start offsets, [40..40+16) cursors
  unsigned int* hist = map->d_counters + 8;
  unsigned int* start = map->d_counters + 24;
  unsigned int* cursor = map->d_counters + 40;
  PCS_CUDA(cudaMemsetAsync(hist, 0, 48 * sizeof(unsigned int), ctx->stream));
  owner_hist_kernel<<<table_grid(ctx, map->n_slots), 256, 0, ctx->stream>>>(map->view, (unsigned int)n_owners, hist);
  count_launch(ctx);
  PCS_TRY(check_launch("owner_hist_kernel"));
  void* h = nullptr;
  PCS_TRY(ctx_pinned(ctx, 256, &h));
  unsigned int* hh = (unsigned int*)h + 16;
  PCS_CUDA(cudaMemcpyAsync(hh, hist, 16 * sizeof(unsigned int), cudaMemcpyDeviceToHost, ctx->stream));
  PCS_CUDA(cudaStreamSynchronize(ctx->stream));
  unsigned int st[16];
  uint64_t total = 0;
  for (int o = 0; o < n_owners; o++)
  {
    st[o] = (unsigned int)total;
    owner_counts[o] = hh[o];
    total += hh[o];
  }
  unsigned int c[2];
  PCS_TRY(read_counters(ctx, map, c));
  if (total > cap_records)
  {
    set_error("records buffer holds %llu, map has %llu voxels", (unsigned long long)cap_records, (unsigned long long)total);
    return PCS_ERR_CAPACITY;
  }
  if (total == 0)
    return PCS_OK;
  PCS_CHECK_ARG(records_dev != nullptr, "records is NULL");
  PCS_CUDA(cudaMemcpyAsync(start, st, (size_t)n_owners * sizeof(unsigned int), cudaMemcpyHostToDevice, ctx->stream));
  owner_scatter_kernel<<<table_grid(ctx, map->n_slots), 256, 0, ctx->stream>>>(map->view, (unsigned int)n_owners, start, cursor,
                                                                             (Record*)records_dev, cap_records);
  count_launch(ctx);
  PCS_TRY(check_launch("owner_scatter_kernel"));
  PCS_CUDA(cudaStreamSynchronize(ctx->stream));  // `st` is a stack buffer
  return PCS_OK;
}

}  // extern "C"
