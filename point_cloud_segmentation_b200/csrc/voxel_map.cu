// voxel_map.cu -- stage 3 state: a device hash of 48-bit voxel keys standing in for the octomap::OcTree that
// OctomapMeshMask::setOctree builds through tesseract_geometry::Octree
// (pcs_scan_integration/src/pcs_scan_integration/octomap_mesh_masking.cpp:105-116).
//
// The path inserts hits only, so a leaf's log-odds is a pure function of its hit count (octomap: +logodds(0.7) per
// hit, clamped at logodds(0.971), early return once clamped) and the pointer tree reduces to a count per key.
// Export re-creates the octree's view: records in Morton (child-index DFS) order with the float log-odds.
#include <cub/device/device_radix_sort.cuh>

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

// ---- peer-memory exchange: every used slot of the local table is added straight into its owner's table ----------------
// The owner tables may live on other GPUs (CUDA IPC mappings reached over NVLink), so the insert uses system-scope
// atomics: a 64-bit CAS claims the slot at the owner's L2, a 64-bit add merges the two counters.
struct OwnerViews
{
  MapView v[16];
};

__device__ __forceinline__ void map_add_system(const MapView& m, unsigned long long key48, unsigned long long add)
{
  const unsigned long long want = key48 | kValidBit;
  unsigned long long slot = slot_hash(key48) & m.mask;
#pragma unroll 1
  for (unsigned int probe = 0; probe < 65536u; ++probe)
  {
    Slot* s = m.slots + slot;
    unsigned long long k = atomicCAS_system(&s->key, 0ull, want);  // one round trip: claims or reads the key
    if (k == 0ull)
    {
      atomicAdd_system(m.n_used, 1u);
      k = want;
    }
    if (k == want)
    {
      atomicAdd_system(&s->counts, add);
      return;
    }
    slot = (slot + 1) & m.mask;
  }
  *m.overflow = 1u;
}

__global__ void map_push_to_owners_kernel(MapView local, OwnerViews owners, unsigned int n_owners)
{
  const unsigned long long n_slots = local.mask + 1;
  for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_slots;
       i += (unsigned long long)gridDim.x * blockDim.x)
  {
    const Slot s = local.slots[i];
    if (!s.key)
      continue;
    const unsigned long long k = s.key & ~kValidBit;
    map_add_system(owners.v[mix64(k) % n_owners], k, s.counts);
  }
}

static unsigned int table_grid(pcs_ctx* ctx, unsigned long long n)
{
  unsigned long long b = (n + 255) / 256;
  const unsigned long long mx = (unsigned long long)ctx->sm_count * 8ull;
  return (unsigned int)(b < mx ? (b ? b : 1) : mx);
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
  cudaError_t e = cudaMalloc(&m->view.slots, n_slots * sizeof(Slot));
  if (e == cudaSuccess)
    e = cudaMalloc(&m->d_counters, 64 * sizeof(unsigned int));
  if (e != cudaSuccess)
  {
    if (m->view.slots)
      cudaFree(m->view.slots);
    delete m;
    return cuda_fail(e, "cudaMalloc(voxel map)", __FILE__, __LINE__);
  }
  m->view.mask = n_slots - 1;
  m->view.n_used = m->d_counters;
  m->view.overflow = m->d_counters + 1;
  m->view.inv_res = 1.0 / resolution;  // octomap: resolution_factor = 1. / resolution
  *out = m;
  return pcs_map_clear(ctx, m);
}

int pcs_map_destroy(pcs_ctx* ctx, pcs_map* map)
{
  if (!map)
    return PCS_OK;
  PCS_TRY(ctx_bind(ctx));
  PCS_CUDA(cudaStreamSynchronize(ctx->stream));
  if (map->is_peer)
  {
    cudaIpcCloseMemHandle(map->view.slots);
    cudaIpcCloseMemHandle(map->d_counters);
  }
  else
  {
    cudaFree(map->view.slots);
    cudaFree(map->d_counters);
  }
  delete map;
  return PCS_OK;
}

int pcs_map_clear(pcs_ctx* ctx, pcs_map* map)
{
  PCS_TRY(ctx_bind(ctx));
  PCS_CHECK_ARG(map != nullptr, "map is NULL");
  PCS_CUDA(cudaMemsetAsync(map->view.slots, 0, map->n_slots * sizeof(Slot), ctx->stream));
  PCS_CUDA(cudaMemsetAsync(map->d_counters, 0, 64 * sizeof(unsigned int), ctx->stream));
  return PCS_OK;
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
  PCS_CUDA(cudaIpcGetMemHandle(&h[0], map->view.slots));
  PCS_CUDA(cudaIpcGetMemHandle(&h[1], map->d_counters));
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
  void *slots = nullptr, *counters = nullptr;
  PCS_CUDA(cudaIpcOpenMemHandle(&slots, h[0], cudaIpcMemLazyEnablePeerAccess));
  cudaError_t e = cudaIpcOpenMemHandle(&counters, h[1], cudaIpcMemLazyEnablePeerAccess);
  if (e != cudaSuccess)
  {
    cudaIpcCloseMemHandle(slots);
    return cuda_fail(e, "cudaIpcOpenMemHandle(counters)", __FILE__, __LINE__);
  }
  pcs_map* m = new (std::nothrow) pcs_map();
  PCS_CHECK_ARG(m != nullptr, "out of host memory");
  m->resolution = resolution;
  m->capacity_voxels = capacity_voxels;
  m->n_slots = n_slots;
  m->is_peer = true;
  m->d_counters = (unsigned int*)counters;
  m->view.slots = (Slot*)slots;
  m->view.mask = n_slots - 1;
  m->view.n_used = m->d_counters;
  m->view.overflow = m->d_counters + 1;
  m->view.inv_res = 1.0 / resolution;
  *out = m;
  return PCS_OK;
}

int pcs_map_push_to_owners_dev(pcs_ctx* ctx, pcs_map* local, pcs_map* const* owners, int n_owners)
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
  map_push_to_owners_kernel<<<table_grid(ctx, local->n_slots), 256, 0, ctx->stream>>>(local->view, ov, (unsigned int)n_owners);
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
