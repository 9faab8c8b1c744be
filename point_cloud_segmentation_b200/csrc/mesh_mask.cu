// mesh_mask.cu -- stage 4: test every mesh triangle against the labelled voxel map and compact the survivors.
//
// Replaces OctomapMeshMask::maskMesh (pcs_scan_integration/src/pcs_scan_integration/octomap_mesh_masking.cpp:118-283):
// the reference builds one Bullet box per occupied octree leaf and one triangle shape per face, runs
// contactTest(ALL) at contact-distance threshold 0.0 (:143,:153) and assembles INSIDE / OUTSIDE / COLORIZED meshes
// from the contacted triangle ids (:163-280).  Here each triangle walks the voxel keys its bounding box covers,
// looks them up in the hash and runs a closed-interval 13-axis separating-axis test in double precision against the
// occupied ones; survivors are emitted in input order by an order-preserving compaction.
//
// Voxel k spans [(k-32768) res, (k-32767) res] per axis, so neighbouring boxes share faces exactly and the result does
// not depend on how the reference's prune() merged octets into larger leaves.
//
// THIS FILE IS COMPILED WITH -fmad=false: the arithmetic must round exactly like oracle/mesh.c (-ffp-contract=off).
//
// Algorithmic bytes per triangle: 16 B face + ~12 B amortised vertices + 1 B flag (+ 88 B per emitted
// triangle); the table lookups hit L2.
#include "common.cuh"
#include "compact.cuh"

namespace pcs
{
constexpr unsigned long long kInlineRangeLimit = 4096;  // voxels a single thread will walk for one triangle

__device__ __forceinline__ double dmin3(double a, double b, double c)
{
  double m = a < b ? a : b;
  return m < c ? m : c;
}
__device__ __forceinline__ double dmax3(double a, double b, double c)
{
  double m = a > b ? a : b;
  return m > c ? m : c;
}

__device__ __forceinline__ bool axis_separates(double ax, double ay, double az, const double* p0, const double* p1,
                                               const double* p2, const double* h)
{
  const double q0 = (ax * p0[0] + ay * p0[1]) + az * p0[2];
  const double q1 = (ax * p1[0] + ay * p1[1]) + az * p1[2];
  const double q2 = (ax * p2[0] + ay * p2[1]) + az * p2[2];
  const double r = (h[0] * fabs(ax) + h[1] * fabs(ay)) + h[2] * fabs(az);
  const double mn = dmin3(q0, q1, q2), mx = dmax3(q0, q1, q2);
  return (mn > r) || (mx < -r);
}

__device__ __forceinline__ bool point_in_box(const double* lo, const double* hi, const double* v)
{
  return v[0] >= lo[0] && v[0] <= hi[0] && v[1] >= lo[1] && v[1] <= hi[1] && v[2] >= lo[2] && v[2] <= hi[2];
}

// closed triangle / axis-aligned box overlap; same operation order as orc_tri_box_overlap
__device__ bool tri_box_overlap(const double* lo, const double* hi, const double* v0, const double* v1, const double* v2)
{
#pragma unroll
  for (int a = 0; a < 3; a++)
  {
    const double mn = dmin3(v0[a], v1[a], v2[a]), mx = dmax3(v0[a], v1[a], v2[a]);
    if (!(mn <= hi[a]) || !(mx >= lo[a]))
      return false;
  }
  // a vertex inside the closed box is a contact outright (exact comparisons, no rounding involved)
  if (point_in_box(lo, hi, v0) || point_in_box(lo, hi, v1) || point_in_box(lo, hi, v2))
    return true;
  double c[3], h[3], p0[3], p1[3], p2[3], e0[3], e1[3], e2[3];
#pragma unroll
  for (int a = 0; a < 3; a++)
  {
    c[a] = 0.5 * (lo[a] + hi[a]);
    h[a] = 0.5 * (hi[a] - lo[a]);
    p0[a] = v0[a] - c[a];
    p1[a] = v1[a] - c[a];
    p2[a] = v2[a] - c[a];
  }
#pragma unroll
  for (int a = 0; a < 3; a++)
  {
    e0[a] = p1[a] - p0[a];
    e1[a] = p2[a] - p1[a];
    e2[a] = p0[a] - p2[a];
  }
  const double* es[3] = { e0, e1, e2 };
#pragma unroll
  for (int j = 0; j < 3; j++)
  {
    const double* e = es[j];
    if (axis_separates(0.0, -e[2], e[1], p0, p1, p2, h))
      return false;
    if (axis_separates(e[2], 0.0, -e[0], p0, p1, p2, h))
      return false;
    if (axis_separates(-e[1], e[0], 0.0, p0, p1, p2, h))
      return false;
  }
  const double nx = e0[1] * e1[2] - e0[2] * e1[1];
  const double ny = e0[2] * e1[0] - e0[0] * e1[2];
  const double nz = e0[0] * e1[1] - e0[1] * e1[0];
  return !axis_separates(nx, ny, nz, p0, p1, p2, h);
}

__device__ __forceinline__ double vox_lo(int k, double res) { return (double)(k - 32768) * res; }
__device__ __forceinline__ double vox_hi(int k, double res) { return (double)(k - 32767) * res; }

// exact interval of keys whose closed extent meets [tmin, tmax]; false if empty / non-finite.
// The interval is clamped to the key range first, so the floor estimate is at most one or two voxels off and the
// correction loops below run a bounded number of steps however far away a (finite) vertex lies.
__device__ bool key_range(double tmin, double tmax, double res, double inv, int& klo, int& khi)
{
  if (!isfinite(tmin) || !isfinite(tmax))
    return false;
  const double range_lo = vox_lo(0, res), range_hi = vox_hi(65535, res);
  if (tmin > range_hi || tmax < range_lo)
    return false;  // entirely outside the map's key range: no voxel can touch it
  tmin = fmax(tmin, range_lo);
  tmax = fmin(tmax, range_hi);
  double e = floor(tmin * inv);
  e = fmin(fmax(e, -40000.0), 40000.0);
  int k = (int)e + 32768;
  while (vox_hi(k - 1, res) >= tmin)
    k--;
  while (vox_hi(k, res) < tmin)
    k++;
  klo = k;
  e = floor(tmax * inv);
  e = fmin(fmax(e, -40000.0), 40000.0);
  k = (int)e + 32768;
  while (vox_lo(k + 1, res) <= tmax)
    k++;
  while (vox_lo(k, res) > tmax)
    k--;
  khi = k;
  klo = max(klo, 0);
  khi = min(khi, 65535);
  return klo <= khi;
}

// is voxel key48 present with count1 >= 1 ?
__device__ __forceinline__ bool map_occupied(const MapView& m, unsigned long long key48)
{
  const unsigned long long want = key48 | kValidBit;
  unsigned long long slot = slot_hash(key48) & m.mask;
  for (unsigned long long probe = 0; probe <= m.mask; ++probe)
  {
    const ulonglong2 s = __ldg(reinterpret_cast<const ulonglong2*>(m.slots + slot));
    if (s.x == want)
      return (s.y >> 32) != 0ull;
    if (s.x == 0ull)
      return false;
    slot = (slot + 1) & m.mask;
  }
  return false;
}

struct TriData
{
  double v0[3], v1[3], v2[3];
  int lo[3], hi[3];
  bool ok;
};

__device__ __forceinline__ void load_tri(const double* __restrict__ verts, unsigned long long nv,
                                         const int* __restrict__ faces4, unsigned long long t, double res, double inv,
                                         TriData& d)
{
  const int4 f = __ldg(reinterpret_cast<const int4*>(faces4) + t);
  d.ok = (unsigned long long)(unsigned int)f.y < nv && (unsigned long long)(unsigned int)f.z < nv &&
         (unsigned long long)(unsigned int)f.w < nv && f.y >= 0 && f.z >= 0 && f.w >= 0;
  if (!d.ok)
    return;
#pragma unroll
  for (int a = 0; a < 3; a++)
  {
    d.v0[a] = __ldg(verts + 3ull * f.y + a);
    d.v1[a] = __ldg(verts + 3ull * f.z + a);
    d.v2[a] = __ldg(verts + 3ull * f.w + a);
  }
#pragma unroll
  for (int a = 0; a < 3; a++)
    d.ok = key_range(dmin3(d.v0[a], d.v1[a], d.v2[a]), dmax3(d.v0[a], d.v1[a], d.v2[a]), res, inv, d.lo[a], d.hi[a]) && d.ok;
}

// Pass 1 (lean, high occupancy): the voxel that holds a vertex decides most triangles of a mesh finer than the voxels
// with one table probe each and no SAT.  Undecided triangles go to a to-do list for pass 2.
__global__ void __launch_bounds__(256, 4) mesh_vertex_probe_kernel(MapView m, double res, const double* __restrict__ verts,
                                                                  unsigned long long nv, const int* __restrict__ faces4,
                                                                  unsigned long long nt, uint8_t* __restrict__ flags,
                                                                  unsigned int* __restrict__ todo,
                                                                  unsigned int* __restrict__ todo_count)
{
  const unsigned long long t = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nt)
    return;
  const int4 f = __ldg(reinterpret_cast<const int4*>(faces4) + t);
  const bool ok = f.y >= 0 && f.z >= 0 && f.w >= 0 && (unsigned long long)f.y < nv && (unsigned long long)f.z < nv &&
                  (unsigned long long)f.w < nv;
  uint8_t flag = 0;
  if (ok)
  {
    const int idx[3] = { f.y, f.z, f.w };
    double v[3][3];
    bool finite = true;
#pragma unroll
    for (int i = 0; i < 3; i++)
#pragma unroll
      for (int a = 0; a < 3; a++)
      {
        v[i][a] = __ldg(verts + 3ull * idx[i] + a);
        finite = finite && isfinite(v[i][a]);
      }
    // one vertex after the other, stopping at the first hit (for a mesh finer than the voxels that is vertex 0)
    for (int i = 0; i < 3 && !flag; i++)
    {
      unsigned long long key = 0;
      bool inside = true;
#pragma unroll
      for (int a = 0; a < 3; a++)
      {
        // e is integral, so e * res and (e + 1) * res are exactly the voxel's bounds (k - 32768) * res, (k - 32767) * res
        const double e = fmin(fmax(floor(v[i][a] * m.inv_res), -32768.0), 32767.0);
        inside = inside && v[i][a] >= e * res && v[i][a] <= (e + 1.0) * res;  // exact test guards the estimate
        key |= (unsigned long long)((int)e + 32768) << (16 * a);
      }
      if (inside && map_occupied(m, key))
        flag = 1;
    }
    if (!finite)
      flag = 0;  // a triangle with a non-finite vertex touches nothing (same rule as the key-range walk)
    else if (!flag)
      todo[atomicAdd(todo_count, 1u)] = (unsigned int)t;
  }
  flags[t] = flag;
}

// Pass 2: triangles the vertex probe could not decide walk the voxel keys their bounding box covers and run the closed
// separating-axis test against the occupied ones.
__global__ void __launch_bounds__(128) mesh_contacts_kernel(MapView m, double res, const double* __restrict__ verts,
                                                           unsigned long long nv, const int* __restrict__ faces4,
                                                           uint8_t* __restrict__ flags, const unsigned int* __restrict__ todo,
                                                           const unsigned int* __restrict__ todo_count,
                                                           unsigned int* __restrict__ big_list,
                                                           unsigned int* __restrict__ big_count)
{
  const unsigned int n_todo = *todo_count;
  for (unsigned int i = blockIdx.x * blockDim.x + threadIdx.x; i < n_todo; i += gridDim.x * blockDim.x)
  {
    const unsigned long long t = todo[i];
    TriData d;
    load_tri(verts, nv, faces4, t, res, m.inv_res, d);
    if (!d.ok)
      continue;
    const unsigned long long vol = (unsigned long long)(d.hi[0] - d.lo[0] + 1) * (unsigned long long)(d.hi[1] - d.lo[1] + 1) *
                                   (unsigned long long)(d.hi[2] - d.lo[2] + 1);
    if (vol > kInlineRangeLimit)
    {
      big_list[atomicAdd(big_count, 1u)] = (unsigned int)t;  // handled voxel-major by mesh_contacts_big_kernel
      continue;
    }
    uint8_t flag = 0;
    for (int kz = d.lo[2]; kz <= d.hi[2] && !flag; kz++)
      for (int ky = d.lo[1]; ky <= d.hi[1] && !flag; ky++)
        for (int kx = d.lo[0]; kx <= d.hi[0] && !flag; kx++)
        {
          const unsigned long long key = ((unsigned long long)kz << 32) | ((unsigned long long)ky << 16) | (unsigned long long)kx;
          if (!map_occupied(m, key))
            continue;
          const double blo[3] = { vox_lo(kx, res), vox_lo(ky, res), vox_lo(kz, res) };
          const double bhi[3] = { vox_hi(kx, res), vox_hi(ky, res), vox_hi(kz, res) };
          if (tri_box_overlap(blo, bhi, d.v0, d.v1, d.v2))
            flag = 1;
        }
    if (flag)
      flags[t] = 1;
  }
}

// triangles whose key range is too large to walk: every occupied voxel tests itself against each of them
__global__ void __launch_bounds__(256) mesh_contacts_big_kernel(MapView m, double res, const double* __restrict__ verts,
                                                               unsigned long long nv, const int* __restrict__ faces4,
                                                               uint8_t* __restrict__ flags,
                                                               const unsigned int* __restrict__ big_list,
                                                               const unsigned int* __restrict__ big_count)
{
  const unsigned int n_big = *big_count;
  if (n_big == 0)
    return;
  const unsigned long long n_slots = m.mask + 1;
  for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_slots;
       i += (unsigned long long)gridDim.x * blockDim.x)
  {
    const Slot s = m.slots[i];
    if (s.key == 0ull || (s.counts >> 32) == 0ull)
      continue;
    const unsigned long long k = s.key & ~kValidBit;
    const int kx = (int)(k & 0xffff), ky = (int)((k >> 16) & 0xffff), kz = (int)((k >> 32) & 0xffff);
    const double blo[3] = { vox_lo(kx, res), vox_lo(ky, res), vox_lo(kz, res) };
    const double bhi[3] = { vox_hi(kx, res), vox_hi(ky, res), vox_hi(kz, res) };
    for (unsigned int b = 0; b < n_big; b++)
    {
      const unsigned int t = big_list[b];
      if (flags[t])
        continue;
      TriData d;
      load_tri(verts, nv, faces4, t, res, m.inv_res, d);
      if (d.ok && tri_box_overlap(blo, bhi, d.v0, d.v1, d.v2))
        flags[t] = 1;
    }
  }
}

// ---- output assembly ------------------------------------------------------------------------------------------------------
struct FlagPred
{
  const uint8_t* flags;
  uint8_t want;
  __device__ bool operator()(unsigned long long t) const { return (flags[t] != 0) == (want != 0); }
};

struct TriEmit
{
  const double* verts;
  const int* faces4;
  double* out_verts;
  int* out_faces4;
  uint8_t* out_rgb;
  unsigned long long cap;
  __device__ void operator()(unsigned long long t, unsigned long long dst) const
  {
    if (dst >= cap)
      return;
    const int4 f = reinterpret_cast<const int4*>(faces4)[t];
    const int idx[3] = { f.y, f.z, f.w };
#pragma unroll
    for (int j = 0; j < 3; j++)
    {
#pragma unroll
      for (int a = 0; a < 3; a++)
        out_verts[9 * dst + 3 * j + a] = verts[3ull * idx[j] + a];
      out_rgb[9 * dst + 3 * j + 0] = 0;  // octomap_mesh_masking.cpp:224,278: all (0,128,0)
      out_rgb[9 * dst + 3 * j + 1] = 128;
      out_rgb[9 * dst + 3 * j + 2] = 0;
    }
    reinterpret_cast<int4*>(out_faces4)[dst] = make_int4(3, (int)(3 * dst), (int)(3 * dst + 1), (int)(3 * dst + 2));
  }
};

struct NullEmit
{
  __device__ void operator()(unsigned long long, unsigned long long) const {}
};

__global__ void fill_rgb_kernel(uint8_t* __restrict__ rgb, unsigned long long nv)
{
  const unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < nv)
  {
    rgb[3 * i] = 0;
    rgb[3 * i + 1] = 128;
    rgb[3 * i + 2] = 0;
  }
}

__global__ void colorize_kernel(const uint8_t* __restrict__ flags, const int* __restrict__ faces4, unsigned long long nt,
                                unsigned long long nv, uint8_t* __restrict__ rgb)
{
  const unsigned long long t = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nt || !flags[t])
    return;
  const int4 f = reinterpret_cast<const int4*>(faces4)[t];
  const int idx[3] = { f.y, f.z, f.w };
#pragma unroll
  for (int j = 0; j < 3; j++)
    if (idx[j] >= 0 && (unsigned long long)idx[j] < nv)
    {
      rgb[3ull * idx[j] + 0] = 255;  // octomap_mesh_masking.cpp:175-180
      rgb[3ull * idx[j] + 1] = 0;
      rgb[3ull * idx[j] + 2] = 0;
    }
}

int mesh_contacts_launch(pcs_ctx* ctx, pcs_map* map, const double* verts_dev, uint64_t nv, const int32_t* faces4_dev,
                         uint64_t nt, uint8_t* flags_dev)
{
  // scratch: [todo_count, big_count (16 B)] [todo list: nt u32] [big list: nt u32]
  void* scr = nullptr;
  PCS_TRY(ctx_scratch(ctx, 16 + 2 * nt * sizeof(unsigned int), &scr));
  unsigned int* todo_count = (unsigned int*)scr;
  unsigned int* big_count = todo_count + 1;
  unsigned int* todo = (unsigned int*)((char*)scr + 16);
  unsigned int* big_list = todo + nt;
  PCS_CUDA(cudaMemsetAsync(scr, 0, 16, ctx->stream));
  const unsigned long long mx = (unsigned long long)ctx->sm_count * 8ull;
  mesh_vertex_probe_kernel<<<(unsigned int)((nt + 255) / 256), 256, 0, ctx->stream>>>(map->view, map->resolution, verts_dev, nv,
                                                                                     faces4_dev, nt, flags_dev, todo, todo_count);
  unsigned long long blocks = (nt + 127) / 128;
  if (blocks > mx)
    blocks = mx;
  mesh_contacts_kernel<<<(unsigned int)blocks, 128, 0, ctx->stream>>>(map->view, map->resolution, verts_dev, nv, faces4_dev,
                                                                    flags_dev, todo, todo_count, big_list, big_count);
  blocks = (map->n_slots + 255) / 256;
  if (blocks > mx)
    blocks = mx;
  mesh_contacts_big_kernel<<<(unsigned int)blocks, 256, 0, ctx->stream>>>(map->view, map->resolution, verts_dev, nv, faces4_dev,
                                                                        flags_dev, big_list, big_count);
  count_launch(ctx, 3);
  return check_launch("mesh_contacts_kernel");
}
}  // namespace pcs

using namespace pcs;

extern "C" {

int pcs_mesh_contacts_dev(pcs_ctx* ctx, pcs_map* map, const double* verts_dev, uint64_t nv, const int32_t* faces4_dev,
                          uint64_t nt, uint8_t* tri_flags_dev)
{
  PCS_TRY(ctx_bind(ctx));
  PCS_CHECK_ARG(map != nullptr, "map is NULL");
  PCS_CHECK_ARG(nt < (1ull << 32) && nv < (1ull << 31), "mesh too large");
  if (nt == 0)
    return PCS_OK;
  PCS_CHECK_ARG(verts_dev && faces4_dev && tri_flags_dev, "NULL argument");
  PCS_CHECK_ARG(((uintptr_t)faces4_dev & 15u) == 0, "faces4 must be 16-byte aligned");
  return mesh_contacts_launch(ctx, map, verts_dev, nv, faces4_dev, nt, tri_flags_dev);
}

int pcs_mesh_mask(pcs_ctx* ctx, pcs_map* map, const double* verts, uint64_t nv, const int32_t* faces4, uint64_t nt,
                  int mask_type, uint64_t out_cap_tris, double* out_verts, int32_t* out_faces4, uint8_t* out_rgb,
                  uint8_t* tri_flags, uint64_t* n_out_verts, uint64_t* n_out_tris)
{
  PCS_TRY(ctx_bind(ctx));
  PCS_CHECK_ARG(map != nullptr, "octree / map not set");
  PCS_CHECK_ARG(n_out_verts && n_out_tris, "NULL count outputs");
  PCS_CHECK_ARG(mask_type >= 0 && mask_type <= 2, "Invalid mask type");
  PCS_CHECK_ARG(nt < (1ull << 31) && nv < (1ull << 31), "mesh too large");
  PCS_CHECK_ARG((nt == 0 || faces4) && (nv == 0 || verts), "input mesh not set");
  *n_out_verts = 0;
  *n_out_tris = 0;
  for (uint64_t t = 0; t < nt; t++)
    PCS_CHECK_ARG(faces4[4 * t] == 3, "faces must be triangles [3, i0, i1, i2]");

  const size_t vb = (size_t)nv * 24, fb = (size_t)nt * 16;
  // staging layout: faces (16-byte aligned) | verts | flags
  void* d_in = nullptr;
  const size_t fb_al = (fb + 255) & ~(size_t)255, vb_al = (vb + 255) & ~(size_t)255;
  PCS_TRY(ctx_staging(ctx, 0, fb_al + vb_al + nt + 256, &d_in));
  int32_t* d_faces = (int32_t*)d_in;
  double* d_verts = (double*)((char*)d_in + fb_al);
  uint8_t* d_flags = (uint8_t*)d_in + fb_al + vb_al;
  if (fb)
    PCS_CUDA(cudaMemcpyAsync(d_faces, faces4, fb, cudaMemcpyHostToDevice, ctx->stream));
  if (vb)
    PCS_CUDA(cudaMemcpyAsync(d_verts, verts, vb, cudaMemcpyHostToDevice, ctx->stream));
  ctx->stats.bytes_h2d += vb + fb;
  if (nt)
    PCS_TRY(mesh_contacts_launch(ctx, map, d_verts, nv, d_faces, nt, d_flags));
  if (tri_flags && nt)
  {
    PCS_CUDA(cudaMemcpyAsync(tri_flags, d_flags, nt, cudaMemcpyDeviceToHost, ctx->stream));
    ctx->stats.bytes_d2h += nt;
  }

  if (mask_type == PCS_RETURN_COLORIZED)
  {
    // octomap_mesh_masking.cpp:163-185: the input mesh itself, coloured per vertex
    *n_out_verts = nv;
    *n_out_tris = nt;
    if (out_verts || out_faces4 || out_rgb)
    {
      if (out_cap_tris < nt)
      {
        set_error("output holds %llu triangles, need %llu", (unsigned long long)out_cap_tris, (unsigned long long)nt);
        return PCS_ERR_CAPACITY;
      }
      if (out_rgb && nv)
      {
        void* d_rgb = nullptr;
        PCS_TRY(ctx_staging(ctx, 1, (size_t)nv * 3, &d_rgb));
        fill_rgb_kernel<<<(unsigned int)((nv + 255) / 256), 256, 0, ctx->stream>>>((uint8_t*)d_rgb, nv);
        if (nt)
          colorize_kernel<<<(unsigned int)((nt + 255) / 256), 256, 0, ctx->stream>>>(d_flags, d_faces, nt, nv, (uint8_t*)d_rgb);
        count_launch(ctx, nt ? 2 : 1);
        PCS_TRY(check_launch("colorize_kernel"));
        PCS_CUDA(cudaMemcpyAsync(out_rgb, d_rgb, (size_t)nv * 3, cudaMemcpyDeviceToHost, ctx->stream));
        ctx->stats.bytes_d2h += (size_t)nv * 3;
      }
      // vertices and faces are returned unchanged: device round trip keeps the data path uniform
      if (out_verts && vb)
        PCS_CUDA(cudaMemcpyAsync(out_verts, d_verts, vb, cudaMemcpyDeviceToHost, ctx->stream));
      if (out_faces4 && fb)
        PCS_CUDA(cudaMemcpyAsync(out_faces4, d_faces, fb, cudaMemcpyDeviceToHost, ctx->stream));
      ctx->stats.bytes_d2h += vb + fb;
    }
    PCS_CUDA(cudaStreamSynchronize(ctx->stream));
    return PCS_OK;
  }

  // INSIDE / OUTSIDE: order-preserving compaction of the triangles with flag == want
  if (nt == 0)
  {
    PCS_CUDA(cudaStreamSynchronize(ctx->stream));
    return PCS_OK;
  }
  const bool emit_now = out_verts && out_faces4 && out_rgb;
  const uint64_t cap = emit_now ? (out_cap_tris < nt ? out_cap_tris : nt) : 0;
  void *scr = nullptr, *d_out = nullptr, *h_total = nullptr;
  PCS_TRY(ctx_pinned(ctx, 64, &h_total));
  const size_t scr_bytes = compact_scratch_bytes(nt);
  // the contact kernels are done with the scratch (stream order), so the compaction can reuse it
  PCS_TRY(ctx_scratch(ctx, scr_bytes, &scr));
  const size_t ov = (size_t)cap * 72, of = (size_t)cap * 16, orgb = (size_t)cap * 9;
  PCS_TRY(ctx_staging(ctx, 1, ((ov + 255) & ~(size_t)255) + ((of + 255) & ~(size_t)255) + orgb + 256, &d_out));
  double* d_ov = (double*)d_out;
  int32_t* d_of = (int32_t*)((char*)d_out + ((ov + 255) & ~(size_t)255));
  uint8_t* d_orgb = (uint8_t*)d_of + ((of + 255) & ~(size_t)255);
  FlagPred pred{ d_flags, (uint8_t)(mask_type == PCS_RETURN_INSIDE ? 1 : 0) };
  const unsigned long long* d_total = nullptr;
  if (emit_now)
  {
    TriEmit emit{ d_verts, d_faces, d_ov, d_of, d_orgb, cap };
    PCS_TRY(compact_run(ctx, pred, emit, nt, scr, &d_total));
  }
  else
    PCS_TRY(compact_run(ctx, pred, NullEmit{}, nt, scr, &d_total));
  PCS_CUDA(cudaMemcpyAsync(h_total, d_total, 8, cudaMemcpyDeviceToHost, ctx->stream));
  PCS_CUDA(cudaStreamSynchronize(ctx->stream));
  const uint64_t kept = *(const unsigned long long*)h_total;
  *n_out_tris = kept;
  *n_out_verts = 3 * kept;
  if (!emit_now)
    return PCS_OK;
  if (kept > cap)
  {
    set_error("output holds %llu triangles, need %llu", (unsigned long long)out_cap_tris, (unsigned long long)kept);
    return PCS_ERR_CAPACITY;
  }
  if (kept)
  {
    PCS_CUDA(cudaMemcpyAsync(out_verts, d_ov, (size_t)kept * 72, cudaMemcpyDeviceToHost, ctx->stream));
    PCS_CUDA(cudaMemcpyAsync(out_faces4, d_of, (size_t)kept * 16, cudaMemcpyDeviceToHost, ctx->stream));
    PCS_CUDA(cudaMemcpyAsync(out_rgb, d_orgb, (size_t)kept * 9, cudaMemcpyDeviceToHost, ctx->stream));
    PCS_CUDA(cudaStreamSynchronize(ctx->stream));
    ctx->stats.bytes_d2h += (size_t)kept * 97;
  }
  return PCS_OK;
}

}  // extern "C"
