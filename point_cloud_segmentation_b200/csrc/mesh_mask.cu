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
__device__ __noinline__ bool tri_box_overlap(const double* lo, const double* hi, const double* v0, const double* v1, const double* v2)
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

// ---- the mesh-mask kernels ---------------------------------------------------------------------------------------------------
//
// Measured on B200 (ncu, round 1 and 2): deciding a triangle is bound by instruction issue (address arithmetic, double
// precision key arithmetic, hash probes), not by HBM traffic or latency.  Most of that work is per VERTEX, and a vertex of
// a welded mesh is shared by ~6 triangles, so it is done once per vertex:
//
//   mesh_vertex_kernel   one thread per vertex: the voxel that holds it (floor estimate, validated with the exact closed
//                        bounds), whether it lies exactly on that voxel's lower / upper faces (then it also touches the
//                        neighbour), and whether that voxel is occupied -> one 8-byte record
//   mesh_tri_kernel      one thread per triangle: the face and its three records.  Any vertex in an occupied voxel is a
//                        contact.  Otherwise the exact key box of the triangle follows from the records alone
//                        (min / max of the vertex keys, widened where a vertex sits on a face -- this equals key_range()
//                        on the coordinates, see vertex_record()); its voxels other than the three vertex voxels (known to
//                        be empty) are probed, and only if one of them is occupied are the nine coordinates fetched for the
//                        closed separating-axis test.  For a mesh finer than the voxels there is usually nothing left to
//                        probe.  Key boxes of 28 .. 4096 voxels are walked by the whole warp, larger ones go to
//                        mesh_contacts_big_kernel (voxel-major).  Triangles with an unusable record (coordinate outside
//                        the key range, estimate not validated) take the coordinate path (key_range()).
constexpr unsigned long long kRecValid = 1ull << 48;      // key bits usable: finite, inside the key range, bounds validated
constexpr unsigned long long kRecNonFinite = 1ull << 49;  // a non-finite coordinate: the triangle touches nothing
constexpr unsigned long long kRecOccupied = 1ull << 50;   // the holding voxel has count1 >= 1
constexpr int kRecOnLowShift = 51;                        // bits 51..53: coordinate == lower bound of the voxel (x, y, z)
constexpr int kRecOnHighShift = 54;                       // bits 54..56: coordinate == upper bound

__device__ __forceinline__ bool finite_d(double x) { return ((unsigned int)__double2hiint(x) & 0x7ff00000u) != 0x7ff00000u; }

// Record of one vertex.  With k validated (vox_lo(k) <= x <= vox_hi(k)) the exact key interval of key_range() for the point
// interval [x, x] is [k - (x == vox_lo(k)), k + (x == vox_hi(k))]: vox_hi(k - 1) and vox_lo(k) are the same expression, so
// the voxel below qualifies iff x equals the shared bound, and likewise above.  key_range() is monotone in its arguments,
// hence a triangle's interval is [min over vertices of the lower ends, max over vertices of the upper ends].
__device__ __forceinline__ unsigned long long vertex_record(const MapView& m, double res, double x, double y, double z)
{
  if (!(finite_d(x) && finite_d(y) && finite_d(z)))
    return kRecNonFinite;
  const double c[3] = { x, y, z };
  unsigned long long rec = kRecValid;
#pragma unroll
  for (int ax = 0; ax < 3; ax++)
  {
    const int k = __double2int_rd(__dmul_rn(m.inv_res, c[ax])) + 32768;  // saturating conversion: far values land outside
    if (k < 0 || k > 65535)
      return 0ull;  // outside the key range: no record, coordinate path
    const double lo = vox_lo(k, res), hi = vox_hi(k, res);
    if (!(c[ax] >= lo && c[ax] <= hi))
      return 0ull;  // the estimate is one voxel off (rounding of inv_res * x): coordinate path
    rec |= (unsigned long long)(unsigned int)k << (16 * ax);
    rec |= (unsigned long long)(c[ax] == lo) << (kRecOnLowShift + ax);
    rec |= (unsigned long long)(c[ax] == hi) << (kRecOnHighShift + ax);
  }
  if (map_occupied(m, rec & 0xffffffffffffull))
    rec |= kRecOccupied;
  return rec;
}

// Programmatic dependent launch: the four kernels of a mask call form a chain in one stream.  Each kernel lets its
// successor be scheduled at once (launch_dependents) and each successor waits (grid_dependency_wait) for its predecessor
// to have completed and flushed before it touches anything the predecessor wrote, so launch latency and the ramp-up of
// one kernel overlap the tail of the one before.
__device__ __forceinline__ void launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void grid_dependency_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

constexpr int kVertPerThread = 2;  // vertices per thread of mesh_vertex_kernel (loads and probes of both in flight)

__global__ void __launch_bounds__(256) mesh_vertex_kernel(MapView m, double res, const double* __restrict__ verts,
                                                         unsigned long long nv, unsigned long long* __restrict__ recs,
                                                         unsigned int* __restrict__ counters)
{
  launch_dependents();
  const unsigned long long i0 = (unsigned long long)blockIdx.x * (blockDim.x * kVertPerThread) + threadIdx.x;
  if (i0 == 0)
    counters[0] = counters[1] = 0u;  // to-do and big-triangle counts of this call (first kernel of the call, stream order)
  double c[kVertPerThread][3];
#pragma unroll
  for (int j = 0; j < kVertPerThread; j++)
  {
    const unsigned long long i = i0 + (unsigned long long)j * blockDim.x;
#pragma unroll
    for (int ax = 0; ax < 3; ax++)
      c[j][ax] = i < nv ? __ldg(verts + 3ull * i + ax) : 0.0;
  }
#pragma unroll
  for (int j = 0; j < kVertPerThread; j++)
  {
    const unsigned long long i = i0 + (unsigned long long)j * blockDim.x;
    const unsigned long long rec = vertex_record(m, res, c[j][0], c[j][1], c[j][2]);
    if (i < nv)
      recs[i] = rec;
  }
}

__device__ __forceinline__ bool face_ok(const int4& f, unsigned long long nv)
{
  return f.y >= 0 && f.z >= 0 && f.w >= 0 && (unsigned long long)f.y < nv && (unsigned long long)f.z < nv &&
         (unsigned long long)f.w < nv;
}

__device__ __forceinline__ bool voxel_hits_tri(double res, int kx, int ky, int kz, const double* v0, const double* v1,
                                               const double* v2)
{
  const double blo[3] = { vox_lo(kx, res), vox_lo(ky, res), vox_lo(kz, res) };
  const double bhi[3] = { vox_hi(kx, res), vox_hi(ky, res), vox_hi(kz, res) };
  return tri_box_overlap(blo, bhi, v0, v1, v2);
}

__device__ __forceinline__ void load_verts(const double* __restrict__ verts, const int4& f, double* v0, double* v1, double* v2)
{
#pragma unroll
  for (int ax = 0; ax < 3; ax++)
  {
    v0[ax] = __ldg(verts + 3ull * f.y + ax);
    v1[ax] = __ldg(verts + 3ull * f.z + ax);
    v2[ax] = __ldg(verts + 3ull * f.w + ax);
  }
}

constexpr int kSmallBoxLimit = 27;  // voxels one thread probes on its own
constexpr int kTriPerThread = 4;    // triangles per thread of mesh_tri_kernel: their loads are issued together

// exact key box of a triangle from its three (valid) records: lower corner and extent per axis
__device__ __forceinline__ void box_from_records(unsigned long long r0, unsigned long long r1, unsigned long long r2, int* lo, int* n)
{
#pragma unroll
  for (int ax = 0; ax < 3; ax++)
  {
    const int k0 = (int)((r0 >> (16 * ax)) & 0xffffull), k1 = (int)((r1 >> (16 * ax)) & 0xffffull),
              k2 = (int)((r2 >> (16 * ax)) & 0xffffull);
    const int l0 = k0 - (int)((r0 >> (kRecOnLowShift + ax)) & 1ull), l1 = k1 - (int)((r1 >> (kRecOnLowShift + ax)) & 1ull),
              l2 = k2 - (int)((r2 >> (kRecOnLowShift + ax)) & 1ull);
    const int h0 = k0 + (int)((r0 >> (kRecOnHighShift + ax)) & 1ull), h1 = k1 + (int)((r1 >> (kRecOnHighShift + ax)) & 1ull),
              h2 = k2 + (int)((r2 >> (kRecOnHighShift + ax)) & 1ull);
    lo[ax] = max(min(min(l0, l1), l2), 0);
    n[ax] = min(max(max(h0, h1), h2), 65535) - lo[ax] + 1;
  }
}

// Common path: one thread takes 4 triangles and issues their 4 face loads, then their 12 record loads, together.
//   * a vertex inside an occupied voxel                                   -> contact
//   * no such vertex, and the key box consists of the vertex voxels only  -> no contact (those voxels are known to be empty;
//     this is the usual case for a mesh finer than the voxels)
//   * anything else (other voxels to probe, unusable record)              -> to-do list of mesh_undecided_kernel, appended
//     with one atomic per warp that has any; the flag is written 0 here and overwritten there on contact.
// The kernel makes no calls and holds no coordinates, so it stays small in registers and code.
__global__ void __launch_bounds__(256, 4) mesh_tri_kernel(const double* __restrict__ verts, unsigned long long nv,
                                                         const int* __restrict__ faces4, unsigned long long nt,
                                                         const unsigned long long* __restrict__ recs,
                                                         uint8_t* __restrict__ flags, unsigned int* __restrict__ todo,
                                                         unsigned int* __restrict__ todo_count)
{
  launch_dependents();
  const unsigned int lane = threadIdx.x & 31u;
  const unsigned long long t0 = (unsigned long long)blockIdx.x * (blockDim.x * kTriPerThread) + threadIdx.x;
  int4 f[kTriPerThread];
#pragma unroll
  for (int j = 0; j < kTriPerThread; j++)
  {
    const unsigned long long t = t0 + (unsigned long long)j * blockDim.x;
    f[j] = make_int4(0, -1, -1, -1);
    if (t < nt)
      f[j] = __ldg(reinterpret_cast<const int4*>(faces4) + t);
  }
  grid_dependency_wait();  // the faces are the caller's; the records come from mesh_vertex_kernel
  unsigned long long r[kTriPerThread][3];
#pragma unroll
  for (int j = 0; j < kTriPerThread; j++)
  {
    r[j][0] = r[j][1] = r[j][2] = kRecNonFinite;  // invalid faces and the padding beyond nt touch nothing
    if (face_ok(f[j], nv))
    {
      r[j][0] = __ldg(recs + f[j].y);
      r[j][1] = __ldg(recs + f[j].z);
      r[j][2] = __ldg(recs + f[j].w);
    }
  }
  unsigned int und_mask = 0;  // bit j: triangle j of this thread goes to the to-do list
#pragma unroll
  for (int j = 0; j < kTriPerThread; j++)
  {
    const unsigned long long t = t0 + (unsigned long long)j * blockDim.x;
    const unsigned long long any = r[j][0] | r[j][1] | r[j][2], all = r[j][0] & r[j][1] & r[j][2];
    const bool finite = !(any & kRecNonFinite);  // a triangle with a non-finite vertex touches nothing
    const bool hit = finite && (all & kRecValid) && (any & kRecOccupied);  // a vertex inside an occupied voxel
    bool undecided = finite && !hit;
    if (undecided && (all & kRecValid))
    {
      int lo[3], n[3];
      box_from_records(r[j][0], r[j][1], r[j][2], lo, n);
      const unsigned long long kmask = 0xffffffffffffull;
      const unsigned long long k0 = r[j][0] & kmask, k1 = r[j][1] & kmask, k2 = r[j][2] & kmask;
      const int distinct = 1 + (k1 != k0 ? 1 : 0) + ((k2 != k0 && k2 != k1) ? 1 : 0);
      // the vertex voxels lie inside the box, so the box holds nothing else iff its volume equals their number
      if (n[0] * n[1] * n[2] == distinct && n[0] <= 3 && n[1] <= 3 && n[2] <= 3)
        undecided = false;
    }
    und_mask |= undecided ? (1u << j) : 0u;
    if (t < nt)
      flags[t] = hit ? 1 : 0;
  }
  // one to-do slot range per warp (a warp with nothing undecided, the usual case, issues no atomic and meets no barrier)
  const unsigned int mine = (unsigned int)__popc(und_mask);
  unsigned int incl = mine;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1)
  {
    const unsigned int v = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= (unsigned int)o)
      incl += v;
  }
  const unsigned int total = __shfl_sync(0xffffffffu, incl, 31);
  if (total == 0)
    return;
  unsigned int base = 0;
  if (lane == 0)
    base = atomicAdd(todo_count, total);
  base = __shfl_sync(0xffffffffu, base, 0);
  unsigned int dst = base + incl - mine;
#pragma unroll
  for (int j = 0; j < kTriPerThread; j++)
    if (und_mask & (1u << j))
      todo[dst++] = (unsigned int)(t0 + (unsigned long long)j * blockDim.x);
}

// The undecided triangles (no vertex in an occupied voxel): exact key box from the records (or from the coordinates
// when a record is unusable), probes of the box voxels other than the vertex voxels, closed separating-axis test where
// one is occupied.  A warp takes 32 consecutive to-do entries so that key boxes of 28 .. 4096 voxels can be walked by the
// whole warp; larger ones go to the big list.
__global__ void __launch_bounds__(128) mesh_undecided_kernel(const MapView m, double res, const double* __restrict__ verts,
                                                            const int* __restrict__ faces4,
                                                            const unsigned long long* __restrict__ recs,
                                                            uint8_t* __restrict__ flags, const unsigned int* __restrict__ todo,
                                                            const unsigned int* __restrict__ todo_count,
                                                            unsigned int* __restrict__ big_list,
                                                            unsigned int* __restrict__ big_count)
{
  launch_dependents();
  grid_dependency_wait();
  const unsigned int n_todo = *todo_count;
  const unsigned int lane = threadIdx.x & 31u;
  const unsigned int n_round = (n_todo + 31u) & ~31u;  // whole warps stay converged for the ballots
  for (unsigned int i = blockIdx.x * blockDim.x + threadIdx.x; i < n_round; i += gridDim.x * blockDim.x)
  {
    const bool live = i < n_todo;
    unsigned int t = 0;
    int4 f = make_int4(3, 0, 0, 0);
    unsigned long long r0 = 0ull, r1 = 0ull, r2 = 0ull;
    if (live)
    {
      t = todo[i];
      f = __ldg(reinterpret_cast<const int4*>(faces4) + t);
      r0 = __ldg(recs + f.y);
      r1 = __ldg(recs + f.z);
      r2 = __ldg(recs + f.w);
    }
    const unsigned long long all = r0 & r1 & r2;
    int lo[3] = { 0, 0, 0 }, n[3] = { 0, 0, 0 };
    bool range_ok = live, have_verts = false, medium = false;
    uint8_t flag = 0;
    double v0[3], v1[3], v2[3];
    if (live)
    {
      if (all & kRecValid)
      {
        box_from_records(r0, r1, r2, lo, n);  // exact key box from the records
      }
      else
      {
        // coordinate path: a vertex outside the key range or with an unvalidated estimate
        load_verts(verts, f, v0, v1, v2);
        have_verts = true;
#pragma unroll
        for (int ax = 0; ax < 3; ax++)
        {
          int hi;
          range_ok = key_range(dmin3(v0[ax], v1[ax], v2[ax]), dmax3(v0[ax], v1[ax], v2[ax]), res, m.inv_res, lo[ax], hi) && range_ok;
          n[ax] = hi - lo[ax] + 1;
        }
      }
    }
    if (range_ok)
    {
      const unsigned long long vol = (unsigned long long)n[0] * (unsigned long long)n[1] * (unsigned long long)n[2];
      if (vol > kInlineRangeLimit)
        big_list[atomicAdd(big_count, 1u)] = t;  // handled voxel-major by mesh_contacts_big_kernel
      else if (vol > (unsigned long long)kSmallBoxLimit)
        medium = true;
      else
      {
        // the vertex voxels are known to be empty when all three records are valid: only the other voxels are probed
        const unsigned long long kmask = 0xffffffffffffull;
        const bool skip = (all & kRecValid) != 0ull;
        // The probes below depend on each other only through the early exit, but each is a round trip to L2 or HBM:
        // touch every home slot (and the coordinates the separating-axis test may need) first, so that the walk finds
        // them in cache.
        for (int dz = 0; dz < n[2]; dz++)
          for (int dy = 0; dy < n[1]; dy++)
            for (int dx = 0; dx < n[0]; dx++)
            {
              const unsigned long long key = ((unsigned long long)(lo[2] + dz) << 32) | ((unsigned long long)(lo[1] + dy) << 16) |
                                             (unsigned long long)(lo[0] + dx);
              if (skip && (key == (r0 & kmask) || key == (r1 & kmask) || key == (r2 & kmask)))
                continue;
              asm volatile("prefetch.global.L2 [%0];" ::"l"(m.slots + (slot_hash(key) & m.mask)));
            }
        if (!have_verts)
        {
          asm volatile("prefetch.global.L2 [%0];" ::"l"(verts + 3ull * f.y));
          asm volatile("prefetch.global.L2 [%0];" ::"l"(verts + 3ull * f.z));
          asm volatile("prefetch.global.L2 [%0];" ::"l"(verts + 3ull * f.w));
        }
        for (int dz = 0; dz < n[2] && !flag; dz++)
          for (int dy = 0; dy < n[1] && !flag; dy++)
            for (int dx = 0; dx < n[0] && !flag; dx++)
            {
              const int kx = lo[0] + dx, ky = lo[1] + dy, kz = lo[2] + dz;
              const unsigned long long key = ((unsigned long long)kz << 32) | ((unsigned long long)ky << 16) | (unsigned long long)kx;
              if (skip && (key == (r0 & kmask) || key == (r1 & kmask) || key == (r2 & kmask)))
                continue;
              if (!map_occupied(m, key))
                continue;
              if (!have_verts)
              {
                load_verts(verts, f, v0, v1, v2);
                have_verts = true;
              }
              if (voxel_hits_tri(res, kx, ky, kz, v0, v1, v2))
                flag = 1;
            }
      }
    }
    // key boxes of 28 .. 4096 voxels: the warp walks them together, 32 voxels per step
    unsigned int walk = __ballot_sync(0xffffffffu, medium);
    while (walk)
    {
      const int src = __ffs(walk) - 1;
      walk &= walk - 1;
      int4 fs;
      fs.x = 3;
      fs.y = __shfl_sync(0xffffffffu, f.y, src);
      fs.z = __shfl_sync(0xffffffffu, f.z, src);
      fs.w = __shfl_sync(0xffffffffu, f.w, src);
      const int bx = __shfl_sync(0xffffffffu, lo[0], src), by = __shfl_sync(0xffffffffu, lo[1], src),
                bz = __shfl_sync(0xffffffffu, lo[2], src);
      const int nx = __shfl_sync(0xffffffffu, n[0], src), ny = __shfl_sync(0xffffffffu, n[1], src),
                nz = __shfl_sync(0xffffffffu, n[2], src);
      double w0[3], w1[3], w2[3];
      load_verts(verts, fs, w0, w1, w2);
      const unsigned int vol = (unsigned int)(nx * ny * nz);
      bool hit = false;
      for (unsigned int base = 0; base < vol && !hit; base += 32)
      {
        const unsigned int q = base + lane;
        bool mine = false;
        if (q < vol)
        {
          const int kx = bx + (int)(q % (unsigned int)nx), ky = by + (int)((q / (unsigned int)nx) % (unsigned int)ny),
                    kz = bz + (int)(q / (unsigned int)(nx * ny));
          const unsigned long long key48 = ((unsigned long long)kz << 32) | ((unsigned long long)ky << 16) | (unsigned long long)kx;
          mine = map_occupied(m, key48) && voxel_hits_tri(res, kx, ky, kz, w0, w1, w2);
        }
        hit = __any_sync(0xffffffffu, mine);
      }
      if ((int)lane == src && hit)
        flag = 1;
    }
    if (live && flag)
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
  grid_dependency_wait();
  const unsigned int n_big = *big_count;
  if (n_big != 0)
  {
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

// Scatter of the kept triangles with coalesced stores.  The generic compact_scatter_kernel would have every kept thread
// write its own 72 + 16 + 9 bytes (8-byte pieces 72 bytes apart across the warp).  Here the kept lanes of a warp -- whose
// outputs are contiguous, the compaction being order preserving -- first lay their nine coordinates down in shared memory
// in output order; the whole warp then streams that run out with consecutive 8-byte words per lane, likewise the colour
// bytes; the 16-byte faces are contiguous per kept lane already.
__global__ void __launch_bounds__(kCompactThreads) tri_emit_scatter_kernel(FlagPred pred, TriEmit emit, unsigned long long n,
                                                                          const unsigned long long* __restrict__ offsets)
{
  __shared__ unsigned int s_warp[kCompactThreads / 32];
  __shared__ double s_v[kCompactThreads / 32][32 * 9];
  const unsigned long long base = (unsigned long long)blockIdx.x * kCompactItems;
  unsigned long long running = offsets[blockIdx.x];
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int r = 0; r < kCompactRounds; r++)
  {
    const unsigned long long t = base + (unsigned long long)r * kCompactThreads + threadIdx.x;
    const bool keep = (t < n) && pred(t);
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
    const unsigned int k = __popc(ballot), rank = __popc(ballot & ((1u << lane) - 1u));
    const unsigned long long first = running + before;  // output index of this warp's first kept triangle
    // room left in the caller's buffers (emit.cap triangles): the count is still exact, the excess is not written
    const unsigned int room = first >= emit.cap ? 0u : (unsigned int)(emit.cap - first < k ? emit.cap - first : k);
    if (keep && rank < room)
    {
      const int4 f = reinterpret_cast<const int4*>(emit.faces4)[t];
      const int idx[3] = { f.y, f.z, f.w };
#pragma unroll
      for (int j = 0; j < 3; j++)
#pragma unroll
        for (int ax = 0; ax < 3; ax++)
          s_v[warp][rank * 9 + 3 * j + ax] = emit.verts[3ull * idx[j] + ax];
      const unsigned long long dst = first + rank;
      reinterpret_cast<int4*>(emit.out_faces4)[dst] = make_int4(3, (int)(3 * dst), (int)(3 * dst + 1), (int)(3 * dst + 2));
    }
    __syncwarp();
    double* ov = emit.out_verts + 9ull * first;
    uint8_t* orgb = emit.out_rgb + 9ull * first;
    for (unsigned int i = lane; i < 9u * room; i += 32)
    {
      ov[i] = s_v[warp][i];
      orgb[i] = (i % 3u == 1u) ? 128 : 0;  // octomap_mesh_masking.cpp:224,278: all (0,128,0)
    }
    running += total;
    __syncthreads();
  }
}

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

template <class... Params, class... Args>
static cudaError_t launch_chained(void (*kernel)(Params...), unsigned int grid, unsigned int block, cudaStream_t stream,
                                  bool dependent, Args... args)
{
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(block);
  cfg.dynamicSmemBytes = 0;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = dependent ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<Params>(args)...);
}

int mesh_contacts_launch(pcs_ctx* ctx, pcs_map* map, const double* verts_dev, uint64_t nv, const int32_t* faces4_dev,
                         uint64_t nt, uint8_t* flags_dev)
{
  // Two counters ([0] to-do, [1] big triangles) in a small context-owned buffer; mesh_vertex_kernel, the first kernel of a
  // call, zeroes them, so no memset is launched.  The to-do list, the big list (nt u32 each) and the vertex records
  // (nv u64) are ordinary scratch.
  if (!ctx->mesh_counters)
    PCS_CUDA(cudaMalloc(&ctx->mesh_counters, 64));
  auto al = [](size_t b) { return (b + 255) & ~(size_t)255; };
  const size_t list_bytes = al(nt * sizeof(unsigned int));
  void* scr = nullptr;
  PCS_TRY(ctx_scratch(ctx, 2 * list_bytes + nv * sizeof(unsigned long long) + 16, &scr));
  unsigned int* counters = (unsigned int*)ctx->mesh_counters;
  unsigned int* todo = (unsigned int*)scr;
  unsigned int* big_list = (unsigned int*)((char*)scr + list_bytes);
  unsigned long long* recs = (unsigned long long*)((char*)scr + 2 * list_bytes);
  const unsigned int mx = (unsigned int)ctx->sm_count * 8u;

  const unsigned int g_vert = (unsigned int)((nv + 256 * kVertPerThread - 1) / (256 * kVertPerThread)) + (nv == 0 ? 1u : 0u);
  const unsigned int g_tri = (unsigned int)((nt + 256 * kTriPerThread - 1) / (256 * kTriPerThread));
  unsigned long long g_und = (nt + 127) / 128;
  if (g_und > mx)
    g_und = mx;
  unsigned long long g_big = (map->n_slots + 255) / 256;
  if (g_big > (unsigned long long)ctx->sm_count)
    g_big = (unsigned long long)ctx->sm_count;  // one CTA per SM: with no big triangle (the usual case) it only starts and exits
  cudaStream_t st = ctx->stream;
  PCS_CUDA(launch_chained(mesh_vertex_kernel, g_vert, 256, st, false, map->view, map->resolution, verts_dev,
                          (unsigned long long)nv, recs, counters));
  PCS_CUDA(launch_chained(mesh_tri_kernel, g_tri, 256, st, true, verts_dev, (unsigned long long)nv, (const int*)faces4_dev,
                          (unsigned long long)nt, (const unsigned long long*)recs, flags_dev, todo, counters));
  PCS_CUDA(launch_chained(mesh_undecided_kernel, (unsigned int)g_und, 128, st, true, map->view, map->resolution, verts_dev,
                          (const int*)faces4_dev, (const unsigned long long*)recs, flags_dev, (const unsigned int*)todo,
                          (const unsigned int*)counters, big_list, counters + 1));
  PCS_CUDA(launch_chained(mesh_contacts_big_kernel, (unsigned int)g_big, 256, st, true, map->view, map->resolution, verts_dev,
                          (unsigned long long)nv, (const int*)faces4_dev, flags_dev, (const unsigned int*)big_list,
                          (const unsigned int*)(counters + 1)));
  count_launch(ctx, 4);
  return check_launch("mesh_tri_kernel");
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

// device-resident mask assembly shared by pcs_mesh_mask and pcs_mesh_mask_dev: flags -> compacted INSIDE / OUTSIDE
// triangles, or per-vertex colours.  *d_total_out points at the device-side count of kept triangles (compaction modes).
static int mesh_assemble_dev(pcs_ctx* ctx, const double* d_verts, uint64_t nv, const int32_t* d_faces, uint64_t nt,
                             const uint8_t* d_flags, int mask_type, uint64_t cap, double* d_ov, int32_t* d_of, uint8_t* d_orgb,
                             const unsigned long long** d_total_out)
{
  *d_total_out = nullptr;
  if (mask_type == PCS_RETURN_COLORIZED)
  {
    if (d_orgb && nv)
    {
      fill_rgb_kernel<<<(unsigned int)((nv + 255) / 256), 256, 0, ctx->stream>>>(d_orgb, nv);
      if (nt)
        colorize_kernel<<<(unsigned int)((nt + 255) / 256), 256, 0, ctx->stream>>>(d_flags, d_faces, nt, nv, d_orgb);
      count_launch(ctx, nt ? 2 : 1);
      PCS_TRY(check_launch("colorize_kernel"));
    }
    return PCS_OK;
  }
  if (nt == 0)
    return PCS_OK;
  void* scr = nullptr;
  // the contact kernels are done with the scratch (stream order), so the compaction can reuse it
  PCS_TRY(ctx_scratch(ctx, compact_scratch_bytes(nt), &scr));
  FlagPred pred{ d_flags, (uint8_t)(mask_type == PCS_RETURN_INSIDE ? 1 : 0) };
  if (d_ov && d_of && d_orgb)
  {
    TriEmit emit{ d_verts, d_faces, d_ov, d_of, d_orgb, cap };
    unsigned long long* offsets = nullptr;
    PCS_TRY(compact_offsets(ctx, pred, nt, scr, &offsets));
    const unsigned int nb = compact_blocks(nt);
    tri_emit_scatter_kernel<<<nb, kCompactThreads, 0, ctx->stream>>>(pred, emit, nt, offsets);
    count_launch(ctx);
    PCS_TRY(check_launch("tri_emit_scatter_kernel"));
    *d_total_out = offsets + nb;
  }
  else
    PCS_TRY(compact_run(ctx, pred, NullEmit{}, nt, scr, d_total_out));
  return PCS_OK;
}

int pcs_mesh_emit_dev(pcs_ctx* ctx, const double* verts_dev, uint64_t nv, const int32_t* faces4_dev, uint64_t nt,
                      const uint8_t* tri_flags_dev, int mask_type, uint64_t out_cap_tris, double* out_verts_dev,
                      int32_t* out_faces4_dev, uint8_t* out_rgb_dev, uint64_t* n_out_tris_dev)
{
  PCS_TRY(ctx_bind(ctx));
  PCS_CHECK_ARG(mask_type >= 0 && mask_type <= 2, "Invalid mask type");
  PCS_CHECK_ARG(nt < (1ull << 31) && nv < (1ull << 31), "mesh too large");
  PCS_CHECK_ARG(nt == 0 || (verts_dev && faces4_dev && tri_flags_dev), "NULL argument");
  PCS_CHECK_ARG(((uintptr_t)faces4_dev & 15u) == 0 && ((uintptr_t)out_faces4_dev & 15u) == 0, "faces4 must be 16-byte aligned");
  PCS_CHECK_ARG(n_out_tris_dev != nullptr && ((uintptr_t)n_out_tris_dev & 7u) == 0, "n_out_tris_dev must be an aligned device pointer");
  const unsigned long long* d_total = nullptr;
  PCS_TRY(mesh_assemble_dev(ctx, verts_dev, nv, faces4_dev, nt, tri_flags_dev, mask_type, out_cap_tris, out_verts_dev,
                            out_faces4_dev, out_rgb_dev, &d_total));
  if (d_total)
    PCS_CUDA(cudaMemcpyAsync(n_out_tris_dev, d_total, 8, cudaMemcpyDeviceToDevice, ctx->stream));
  else
    PCS_CUDA(cudaMemsetAsync(n_out_tris_dev, 0, 8, ctx->stream));  // COLORIZED (all nt kept) and empty meshes: no compaction ran
  return PCS_OK;
}

int pcs_mesh_mask_dev(pcs_ctx* ctx, pcs_map* map, const double* verts_dev, uint64_t nv, const int32_t* faces4_dev, uint64_t nt,
                      int mask_type, uint64_t out_cap_tris, double* out_verts_dev, int32_t* out_faces4_dev,
                      uint8_t* out_rgb_dev, uint8_t* tri_flags_dev, uint64_t* n_out_tris_dev)
{
  PCS_TRY(ctx_bind(ctx));
  PCS_CHECK_ARG(map != nullptr, "octree / map not set");
  PCS_CHECK_ARG(mask_type >= 0 && mask_type <= 2, "Invalid mask type");
  PCS_CHECK_ARG(nt < (1ull << 31) && nv < (1ull << 31), "mesh too large");
  PCS_CHECK_ARG(tri_flags_dev != nullptr || nt == 0, "tri_flags_dev is NULL");
  PCS_CHECK_ARG(((uintptr_t)faces4_dev & 15u) == 0 && ((uintptr_t)out_faces4_dev & 15u) == 0, "faces4 must be 16-byte aligned");
  PCS_CHECK_ARG(((uintptr_t)n_out_tris_dev & 7u) == 0, "n_out_tris_dev must be 8-byte aligned");
  if (nt)
  {
    PCS_CHECK_ARG(verts_dev && faces4_dev, "input mesh not set");
    PCS_TRY(mesh_contacts_launch(ctx, map, verts_dev, nv, faces4_dev, nt, tri_flags_dev));
  }
  const unsigned long long* d_total = nullptr;
  PCS_TRY(mesh_assemble_dev(ctx, verts_dev, nv, faces4_dev, nt, tri_flags_dev, mask_type, out_cap_tris, out_verts_dev,
                            out_faces4_dev, out_rgb_dev, &d_total));
  if (n_out_tris_dev)
  {
    if (d_total)
      PCS_CUDA(cudaMemcpyAsync(n_out_tris_dev, d_total, 8, cudaMemcpyDeviceToDevice, ctx->stream));
    else
    {
      // COLORIZED keeps every triangle; an empty mesh keeps none
      void* h = nullptr;
      PCS_TRY(ctx_pinned(ctx, 64, &h));
      PCS_CUDA(cudaStreamSynchronize(ctx->stream));  // the pinned word may still be in use by an earlier copy
      *(unsigned long long*)((char*)h + 32) = (mask_type == PCS_RETURN_COLORIZED) ? nt : 0ull;
      PCS_CUDA(cudaMemcpyAsync(n_out_tris_dev, (char*)h + 32, 8, cudaMemcpyHostToDevice, ctx->stream));
      PCS_CUDA(cudaStreamSynchronize(ctx->stream));
    }
  }
  return PCS_OK;
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
        const unsigned long long* unused = nullptr;
        PCS_TRY(mesh_assemble_dev(ctx, d_verts, nv, d_faces, nt, d_flags, mask_type, nt, nullptr, nullptr, (uint8_t*)d_rgb, &unused));
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
  void *d_out = nullptr, *h_total = nullptr;
  PCS_TRY(ctx_pinned(ctx, 64, &h_total));
  const size_t ov = (size_t)cap * 72, of = (size_t)cap * 16, orgb = (size_t)cap * 9;
  PCS_TRY(ctx_staging(ctx, 1, ((ov + 255) & ~(size_t)255) + ((of + 255) & ~(size_t)255) + orgb + 256, &d_out));
  double* d_ov = (double*)d_out;
  int32_t* d_of = (int32_t*)((char*)d_out + ((ov + 255) & ~(size_t)255));
  uint8_t* d_orgb = (uint8_t*)d_of + ((of + 255) & ~(size_t)255);
  const unsigned long long* d_total = nullptr;
  PCS_TRY(mesh_assemble_dev(ctx, d_verts, nv, d_faces, nt, d_flags, mask_type, cap, emit_now ? d_ov : nullptr,
                            emit_now ? d_of : nullptr, emit_now ? d_orgb : nullptr, &d_total));
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
