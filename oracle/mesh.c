/*
 * oracle/mesh.c -- TEST INFRASTRUCTURE (see pcs_oracle.h).  CPU restatement of OctomapMeshMask::maskMesh and the
 * mesh I/O around it.
 *
 *   contact test   pcs_scan_integration/src/pcs_scan_integration/octomap_mesh_masking.cpp:118-155: Bullet discrete
 *                  contactTest(ALL) of one box per occupied octree leaf against one triangle per mesh face at
 *                  contact-distance threshold 0.0, identity poses -> a (box, triangle) pair is reported iff their
 *                  distance is <= 0 (touching or penetrating).  Bullet / tesseract_collision are not in the tree;
 *                  restated as a closed-interval 13-axis separating-axis test on doubles.  Pinned by the
 *                  reference golden OUTSIDE = 11400 triangles (octomap_mesh_masking_unit.cpp:182-183) which this
 *                  reproduces exactly on box_2m.bt x plane_4m.stl; INSIDE (:170-171, 1975 entries) is Bullet's
 *                  per-contact multiset -- geometry gives 1976 pairs over 888 unique triangles.
 *   assembly       octomap_mesh_masking.cpp:163-280 (COLORIZED / INSIDE / OUTSIDE); INSIDE here emits each
 *                  contacted triangle once, in input order (set semantics).
 *   mesh load      octomap_mesh_masking.cpp:98-103 -> tesseract createMeshFromPath -> assimp with
 *                  Triangulate | JoinIdenticalVertices (normals removed): pins 6273 / 8 vertices
 *                  (octomap_mesh_masking_unit.cpp:76-77,85-86).
 *   mesh save      octomap_mesh_masking.cpp:285-292 -> tesseract writeSimplePlyFile (ASCII PLY, per-vertex uchar
 *                  colours, "%.7f" coordinates); restated from upstream, PARITY UNPINNED.
 * Voxel boxes: the leaf with min-corner key k and edge 2^l voxels spans [(k-32768)*res, (k+2^l-32768)*res] per axis,
 * so neighbouring boxes share their faces exactly and a pruned leaf equals the union of its children.
 * Build with -ffp-contract=off.
 */
#include "pcs_oracle.h"

#include <ctype.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

static inline double dmin3(double a, double b, double c)
{
  double m = a < b ? a : b;
  return m < c ? m : c;
}
static inline double dmax3(double a, double b, double c)
{
  double m = a > b ? a : b;
  return m > c ? m : c;
}

/* one cross-product axis a = (ax, ay, az); p* translated vertices; h half extents */
static inline int axis_separates(double ax, double ay, double az, const double* p0, const double* p1, const double* p2,
                                 const double* h)
{
  const double q0 = (ax * p0[0] + ay * p0[1]) + az * p0[2];
  const double q1 = (ax * p1[0] + ay * p1[1]) + az * p1[2];
  const double q2 = (ax * p2[0] + ay * p2[1]) + az * p2[2];
  const double r = (h[0] * fabs(ax) + h[1] * fabs(ay)) + h[2] * fabs(az);
  const double mn = dmin3(q0, q1, q2), mx = dmax3(q0, q1, q2);
  return (mn > r) || (mx < -r);
}

int orc_tri_box_overlap(const double lo[3], const double hi[3], const double v0[3], const double v1[3],
                        const double v2[3])
{
  /* 3 box axes: exact comparisons on the untranslated coordinates */
  for (int a = 0; a < 3; a++)
  {
    const double mn = dmin3(v0[a], v1[a], v2[a]), mx = dmax3(v0[a], v1[a], v2[a]);
    if (mn > hi[a] || mx < lo[a])
      return 0;
    if (!(mn <= hi[a]) || !(mx >= lo[a])) /* NaN vertices never overlap */
      return 0;
  }
  /* a vertex inside the closed box is a contact outright (exact comparisons, no rounding involved) */
  const double* vs[3] = { v0, v1, v2 };
  for (int i = 0; i < 3; i++)
    if (vs[i][0] >= lo[0] && vs[i][0] <= hi[0] && vs[i][1] >= lo[1] && vs[i][1] <= hi[1] && vs[i][2] >= lo[2] &&
        vs[i][2] <= hi[2])
      return 1;
  double c[3], h[3], p0[3], p1[3], p2[3], e0[3], e1[3], e2[3];
  for (int a = 0; a < 3; a++)
  {
    c[a] = 0.5 * (lo[a] + hi[a]);
    h[a] = 0.5 * (hi[a] - lo[a]);
    p0[a] = v0[a] - c[a];
    p1[a] = v1[a] - c[a];
    p2[a] = v2[a] - c[a];
  }
  for (int a = 0; a < 3; a++)
  {
    e0[a] = p1[a] - p0[a];
    e1[a] = p2[a] - p1[a];
    e2[a] = p0[a] - p2[a];
  }
  /* 9 axes: unit box axis u_i x e_j.  u_x x e = (0, -e.z, e.y); u_y x e = (e.z, 0, -e.x); u_z x e = (-e.y, e.x, 0) */
  const double* es[3] = { e0, e1, e2 };
  for (int j = 0; j < 3; j++)
  {
    const double* e = es[j];
    if (axis_separates(0.0, -e[2], e[1], p0, p1, p2, h))
      return 0;
    if (axis_separates(e[2], 0.0, -e[0], p0, p1, p2, h))
      return 0;
    if (axis_separates(-e[1], e[0], 0.0, p0, p1, p2, h))
      return 0;
  }
  /* triangle normal n = e0 x e1 */
  const double nx = e0[1] * e1[2] - e0[2] * e1[1];
  const double ny = e0[2] * e1[0] - e0[0] * e1[2];
  const double nz = e0[0] * e1[1] - e0[1] * e1[0];
  if (axis_separates(nx, ny, nz, p0, p1, p2, h))
    return 0;
  return 1;
}

static inline void tri_verts(const double* verts, const int32_t* faces4, size_t t, const double** a, const double** b,
                             const double** c)
{
  *a = verts + 3 * (size_t)faces4[4 * t + 1];
  *b = verts + 3 * (size_t)faces4[4 * t + 2];
  *c = verts + 3 * (size_t)faces4[4 * t + 3];
}

uint64_t orc_mesh_contacts_brute(const double* box_lo, const double* box_hi, size_t nb, const double* verts,
                                 const int32_t* faces4, size_t nt, uint32_t* pair_count)
{
  uint64_t total = 0;
  for (size_t t = 0; t < nt; t++)
  {
    const double *a, *b, *c;
    tri_verts(verts, faces4, t, &a, &b, &c);
    uint32_t n = 0;
    for (size_t i = 0; i < nb; i++)
      n += (uint32_t)orc_tri_box_overlap(box_lo + 3 * i, box_hi + 3 * i, a, b, c);
    pair_count[t] = n;
    total += n;
  }
  return total;
}

/* ---- against the pointer octree (boxes = occupied leaves at their own depth) ------------------------ */
typedef struct
{
  const orc_octree* t;
  double res;
  const double *a, *b, *c;
  double tmin[3], tmax[3];
  uint32_t count;
} tq;

/* walk the tree through the public search API would cost a descent per voxel; use the leaf list instead, bucketed
 * once by an implicit recursive subdivision: the leaves come out in DFS order so a stack of (start,end) suffices.
 * Simpler and fast enough for an oracle: recursive descent over key space with orc_octree_search at the leaves of
 * the recursion. */
static void contacts_recurs(tq* q, uint32_t x, uint32_t y, uint32_t z, int depth)
{
  const uint32_t size = 1u << (16 - depth);
  const double lo[3] = { (double)((int64_t)x - 32768) * q->res, (double)((int64_t)y - 32768) * q->res,
                         (double)((int64_t)z - 32768) * q->res };
  const double hi[3] = { (double)((int64_t)x + size - 32768) * q->res, (double)((int64_t)y + size - 32768) * q->res,
                         (double)((int64_t)z + size - 32768) * q->res };
  for (int a = 0; a < 3; a++)
    if (q->tmin[a] > hi[a] || q->tmax[a] < lo[a])
      return;
  float lv = 0.0f;
  int ld = 0;
  if (!orc_octree_search(q->t, (uint16_t)x, (uint16_t)y, (uint16_t)z, &lv, &ld))
  {
    /* search() returns NULL either when this region is unknown or when the descent left the tree below `depth`;
     * in the second case finer cells may still exist, so keep subdividing until the leaf level. */
    if (depth == 16)
      return;
  }
  else if (ld <= depth)
  {
    /* a leaf that covers this whole cell */
    if (ld == depth && lv >= 0.0f)
      q->count += (uint32_t)orc_tri_box_overlap(lo, hi, q->a, q->b, q->c);
    return;
  }
  if (depth == 16)
    return;
  const uint32_t half = size >> 1;
  for (int i = 0; i < 8; i++)
    contacts_recurs(q, x + ((i & 1) ? half : 0), y + ((i & 2) ? half : 0), z + ((i & 4) ? half : 0), depth + 1);
}

uint64_t orc_mesh_contacts_tree(const orc_octree* t, const double* verts, size_t nv, const int32_t* faces4, size_t nt,
                                uint32_t* pair_count)
{
  (void)nv;
  uint64_t total = 0;
  for (size_t i = 0; i < nt; i++)
  {
    tq q;
    q.t = t;
    q.res = orc_octree_resolution(t);
    tri_verts(verts, faces4, i, &q.a, &q.b, &q.c);
    for (int a = 0; a < 3; a++)
    {
      q.tmin[a] = dmin3(q.a[a], q.b[a], q.c[a]);
      q.tmax[a] = dmax3(q.a[a], q.b[a], q.c[a]);
    }
    q.count = 0;
    contacts_recurs(&q, 0, 0, 0, 0);
    pair_count[i] = q.count;
    total += q.count;
  }
  return total;
}

/* ---- against a flat list of occupied finest-level voxels -------------------------------------------- */
static int u64_cmp(const void* a, const void* b)
{
  const uint64_t x = *(const uint64_t*)a, y = *(const uint64_t*)b;
  return x < y ? -1 : (x > y ? 1 : 0);
}

static inline double vox_lo(int k, double res) { return (double)(k - 32768) * res; }
static inline double vox_hi(int k, double res) { return (double)(k - 32767) * res; }

/* exact candidate key interval [klo, khi] of voxels whose closed extent meets [tmin, tmax] on one axis */
static int key_range(double tmin, double tmax, double res, double inv, int* klo, int* khi)
{
  if (!isfinite(tmin) || !isfinite(tmax))
    return 0;
  /* clamp to the key range first: the correction loops then run a bounded number of steps for any finite vertex */
  const double range_lo = vox_lo(0, res), range_hi = vox_hi(65535, res);
  if (tmin > range_hi || tmax < range_lo)
    return 0;
  if (tmin < range_lo)
    tmin = range_lo;
  if (tmax > range_hi)
    tmax = range_hi;
  double e = floor(tmin * inv);
  if (e < -40000.0)
    e = -40000.0;
  if (e > 40000.0)
    e = 40000.0;
  int k = (int)e + 32768;
  while (vox_hi(k - 1, res) >= tmin)
    k--;
  while (vox_hi(k, res) < tmin)
    k++;
  *klo = k;
  e = floor(tmax * inv);
  if (e < -40000.0)
    e = -40000.0;
  if (e > 40000.0)
    e = 40000.0;
  k = (int)e + 32768;
  while (vox_lo(k + 1, res) <= tmax)
    k++;
  while (vox_lo(k, res) > tmax)
    k--;
  *khi = k;
  if (*klo < 0)
    *klo = 0;
  if (*khi > 65535)
    *khi = 65535;
  return *klo <= *khi;
}

uint64_t orc_mesh_contacts_keys(const uint64_t* keys_in, size_t nk, double resolution, const double* verts, size_t nv,
                                const int32_t* faces4, size_t nt, uint8_t* tri_flags)
{
  (void)nv;
  uint64_t* keys = (uint64_t*)malloc(sizeof(uint64_t) * (nk ? nk : 1));
  memcpy(keys, keys_in, sizeof(uint64_t) * nk);
  qsort(keys, nk, sizeof(uint64_t), u64_cmp);
  const double inv = 1.0 / resolution;
  uint64_t hit = 0;
  for (size_t t = 0; t < nt; t++)
  {
    const double *a, *b, *c;
    tri_verts(verts, faces4, t, &a, &b, &c);
    int lo[3], hi[3], ok = 1;
    for (int ax = 0; ax < 3 && ok; ax++)
      ok = key_range(dmin3(a[ax], b[ax], c[ax]), dmax3(a[ax], b[ax], c[ax]), resolution, inv, &lo[ax], &hi[ax]);
    uint8_t flag = 0;
    if (ok)
      for (int kz = lo[2]; kz <= hi[2] && !flag; kz++)
        for (int ky = lo[1]; ky <= hi[1] && !flag; ky++)
          for (int kx = lo[0]; kx <= hi[0] && !flag; kx++)
          {
            const uint64_t key = ((uint64_t)kz << 32) | ((uint64_t)ky << 16) | (uint64_t)kx;
            if (!bsearch(&key, keys, nk, sizeof(uint64_t), u64_cmp))
              continue;
            const double blo[3] = { vox_lo(kx, resolution), vox_lo(ky, resolution), vox_lo(kz, resolution) };
            const double bhi[3] = { vox_hi(kx, resolution), vox_hi(ky, resolution), vox_hi(kz, resolution) };
            if (orc_tri_box_overlap(blo, bhi, a, b, c))
              flag = 1;
          }
    tri_flags[t] = flag;
    hit += flag;
  }
  free(keys);
  return hit;
}

size_t orc_mask_mesh_emit(const double* verts, size_t nv, const int32_t* faces4, size_t nt, const uint8_t* tri_flags,
                          int mask_type, double* out_verts, int32_t* out_faces4, uint8_t* out_rgb, size_t* n_out_verts)
{
  if (mask_type == 2)
  {
    /* octomap_mesh_masking.cpp:163-185 */
    memcpy(out_verts, verts, nv * 3 * sizeof(double));
    memcpy(out_faces4, faces4, nt * 4 * sizeof(int32_t));
    for (size_t i = 0; i < nv; i++)
    {
      out_rgb[3 * i] = 0;
      out_rgb[3 * i + 1] = 128;
      out_rgb[3 * i + 2] = 0;
    }
    for (size_t t = 0; t < nt; t++)
      if (tri_flags[t])
        for (int j = 1; j <= 3; j++)
        {
          const size_t v = (size_t)faces4[4 * t + j];
          out_rgb[3 * v] = 255;
          out_rgb[3 * v + 1] = 0;
          out_rgb[3 * v + 2] = 0;
        }
    *n_out_verts = nv;
    return nt;
  }
  /* :186-226 (INSIDE, one entry per contacted triangle) and :227-280 (OUTSIDE), both in input order, un-welded */
  const int want = (mask_type == 0) ? 1 : 0;
  size_t n = 0;
  for (size_t t = 0; t < nt; t++)
  {
    if ((tri_flags[t] != 0) != want)
      continue;
    for (int j = 0; j < 3; j++)
    {
      const double* v = verts + 3 * (size_t)faces4[4 * t + 1 + j];
      out_verts[9 * n + 3 * j + 0] = v[0];
      out_verts[9 * n + 3 * j + 1] = v[1];
      out_verts[9 * n + 3 * j + 2] = v[2];
      out_rgb[9 * n + 3 * j + 0] = 0;
      out_rgb[9 * n + 3 * j + 1] = 128;
      out_rgb[9 * n + 3 * j + 2] = 0;
    }
    out_faces4[4 * n + 0] = 3;
    out_faces4[4 * n + 1] = (int32_t)(3 * n);
    out_faces4[4 * n + 2] = (int32_t)(3 * n + 1);
    out_faces4[4 * n + 3] = (int32_t)(3 * n + 2);
    n++;
  }
  *n_out_verts = 3 * n;
  return n;
}

/* ---- mesh I/O ------------------------------------------------------------------------------------ */
typedef struct
{
  uint32_t bits[3];
  int32_t index;
  int used;
} wslot;

typedef struct
{
  wslot* slots;
  size_t mask;
} welder;

static void welder_init(welder* w, size_t n)
{
  size_t cap = 16;
  while (cap < 2 * n + 1)
    cap <<= 1;
  w->slots = (wslot*)calloc(cap, sizeof(wslot));
  w->mask = cap - 1;
}

/* returns the index of the vertex equal to (x,y,z) (bitwise, -0 folded to +0) or assigns `next` */
static int32_t welder_get(welder* w, const float v[3], int32_t next, int* is_new)
{
  uint32_t b[3];
  for (int i = 0; i < 3; i++)
  {
    float f = v[i] == 0.0f ? 0.0f : v[i];
    memcpy(&b[i], &f, 4);
  }
  uint64_t hsh = 1469598103934665603ULL;
  for (int i = 0; i < 3; i++)
    hsh = (hsh ^ b[i]) * 1099511628211ULL;
  size_t s = (size_t)(hsh ^ (hsh >> 29)) & w->mask;
  for (;;)
  {
    wslot* sl = &w->slots[s];
    if (!sl->used)
    {
      sl->used = 1;
      memcpy(sl->bits, b, sizeof(b));
      sl->index = next;
      *is_new = 1;
      return next;
    }
    if (sl->bits[0] == b[0] && sl->bits[1] == b[1] && sl->bits[2] == b[2])
    {
      *is_new = 0;
      return sl->index;
    }
    s = (s + 1) & w->mask;
  }
}

int orc_load_stl_binary(const uint8_t* bytes, size_t n, double* verts, size_t vcap, int32_t* faces4, size_t tcap,
                        size_t* nv, size_t* nt)
{
  if (n < 84)
    return -1;
  uint32_t count;
  memcpy(&count, bytes + 80, 4);
  if (n < 84 + (size_t)count * 50)
    return -1;
  welder w;
  welder_init(&w, (size_t)count * 3);
  size_t nvert = 0;
  for (uint32_t t = 0; t < count; t++)
  {
    const uint8_t* rec = bytes + 84 + (size_t)t * 50 + 12; /* skip the facet normal */
    if (t < tcap)
      faces4[4 * (size_t)t] = 3;
    for (int j = 0; j < 3; j++)
    {
      float v[3];
      memcpy(v, rec + 12 * j, 12);
      int is_new = 0;
      const int32_t idx = welder_get(&w, v, (int32_t)nvert, &is_new);
      if (is_new)
      {
        if (nvert < vcap)
        {
          verts[3 * nvert] = v[0];
          verts[3 * nvert + 1] = v[1];
          verts[3 * nvert + 2] = v[2];
        }
        nvert++;
      }
      if (t < tcap)
        faces4[4 * (size_t)t + 1 + j] = idx;
    }
  }
  free(w.slots);
  *nv = nvert;
  *nt = count;
  return (nvert <= vcap && count <= tcap) ? 0 : 1;
}

static const char* next_line(const char* p, const char* end, char* buf, size_t bufsz)
{
  size_t k = 0;
  while (p < end && *p != '\n')
  {
    if (k + 1 < bufsz && *p != '\r')
      buf[k++] = *p;
    p++;
  }
  buf[k] = 0;
  return p < end ? p + 1 : end;
}

int orc_load_ply_ascii(const char* text, size_t n, double* verts, size_t vcap, int32_t* faces4, size_t tcap, size_t* nv,
                       size_t* nt)
{
  const char* p = text;
  const char* end = text + n;
  char line[512];
  size_t n_vertex = 0, n_face = 0;
  int vprops = 0, in_vertex = 0, ascii = 0;
  p = next_line(p, end, line, sizeof(line));
  if (strncmp(line, "ply", 3))
    return -1;
  for (;;)
  {
    if (p >= end)
      return -1;
    p = next_line(p, end, line, sizeof(line));
    if (!strncmp(line, "format ascii", 12))
      ascii = 1;
    else if (!strncmp(line, "element vertex", 14))
    {
      n_vertex = (size_t)atoll(line + 14);
      in_vertex = 1;
    }
    else if (!strncmp(line, "element face", 12))
    {
      n_face = (size_t)atoll(line + 12);
      in_vertex = 0;
    }
    else if (!strncmp(line, "element", 7))
      in_vertex = 0;
    else if (!strncmp(line, "property", 8) && in_vertex)
      vprops++;
    else if (!strncmp(line, "end_header", 10))
      break;
  }
  if (!ascii || vprops < 3)
    return -1;
  welder w;
  welder_init(&w, n_vertex);
  int32_t* remap = (int32_t*)malloc(sizeof(int32_t) * (n_vertex ? n_vertex : 1));
  size_t nvert = 0;
  for (size_t i = 0; i < n_vertex; i++)
  {
    p = next_line(p, end, line, sizeof(line));
    float v[3];
    if (sscanf(line, "%f %f %f", &v[0], &v[1], &v[2]) != 3)
    {
      free(remap);
      free(w.slots);
      return -1;
    }
    int is_new = 0;
    remap[i] = welder_get(&w, v, (int32_t)nvert, &is_new);
    if (is_new)
    {
      if (nvert < vcap)
      {
        verts[3 * nvert] = v[0];
        verts[3 * nvert + 1] = v[1];
        verts[3 * nvert + 2] = v[2];
      }
      nvert++;
    }
  }
  size_t ntri = 0;
  for (size_t i = 0; i < n_face; i++)
  {
    p = next_line(p, end, line, sizeof(line));
    char* q = line;
    const long cnt = strtol(q, &q, 10);
    long idx[64];
    for (long j = 0; j < cnt && j < 64; j++)
      idx[j] = strtol(q, &q, 10);
    for (long j = 1; j + 1 < cnt; j++) /* fan triangulation */
    {
      if (ntri < tcap)
      {
        faces4[4 * ntri] = 3;
        faces4[4 * ntri + 1] = remap[idx[0]];
        faces4[4 * ntri + 2] = remap[idx[j]];
        faces4[4 * ntri + 3] = remap[idx[j + 1]];
      }
      ntri++;
    }
  }
  free(remap);
  free(w.slots);
  *nv = nvert;
  *nt = ntri;
  return (nvert <= vcap && ntri <= tcap) ? 0 : 1;
}

size_t orc_write_ply_ascii(const double* verts, size_t nv, const uint8_t* rgb, const int32_t* faces4, size_t nt, char* out,
                           size_t cap)
{
  size_t n = 0;
#define EMIT(...)                                                                                                      \
  do                                                                                                                   \
  {                                                                                                                    \
    int k_ = snprintf(out ? out + (n < cap ? n : cap) : NULL, out && n < cap ? cap - n : 0, __VA_ARGS__);               \
    n += (size_t)k_;                                                                                                   \
  } while (0)
  EMIT("ply\nformat ascii 1.0\ncomment made by tesseract\nelement vertex %zu\n", nv);
  EMIT("property float x\nproperty float y\nproperty float z\n");
  if (rgb)
    EMIT("property uchar red\nproperty uchar green\nproperty uchar blue\n");
  EMIT("element face %zu\nproperty list uchar int vertex_indices\nend_header\n", nt);
  for (size_t i = 0; i < nv; i++)
  {
    if (rgb)
      EMIT("%.7f %.7f %.7f %d %d %d\n", verts[3 * i], verts[3 * i + 1], verts[3 * i + 2], rgb[3 * i], rgb[3 * i + 1],
           rgb[3 * i + 2]);
    else
      EMIT("%.7f %.7f %.7f\n", verts[3 * i], verts[3 * i + 1], verts[3 * i + 2]);
  }
  for (size_t t = 0; t < nt; t++)
    EMIT("%d %d %d %d\n", faces4[4 * t], faces4[4 * t + 1], faces4[4 * t + 2], faces4[4 * t + 3]);
#undef EMIT
  return n;
}
