/*
 * oracle/octree.c -- TEST INFRASTRUCTURE (see pcs_oracle.h).  Pointer octree restating what
 * OctomapMeshMask::setOctree reaches (pcs_scan_integration/src/pcs_scan_integration/octomap_mesh_masking.cpp:105-116):
 *   tesseract_geometry::Octree(cloud, resolution, BOX, prune=true)
 *     -> for each point octomap::OcTree::updateNode(x, y, z, true)
 *     -> updateInnerOccupancy()
 *     -> tesseract's prune (collapse octets whose 8 children are occupied leaves; parent copies child 0)
 * octomap and tesseract are not in the reference tree (dependencies.rosinstall:7-17, unpinned branches); the rules
 * below restate octomap 1.9's OccupancyOcTreeBase / OcTreeBaseImpl from their published sources:
 *   key      k = (int)floor((1/res) * coord) + 32768, valid iff 0 <= k < 65536   (coordToKeyChecked)
 *   update   early return when the found leaf already sits at the clamp; recursive descent creating children or
 *            expanding a pruned node; leaf += logodds(0.7) clamped to [logodds(0.1192), logodds(0.971)]; on the way
 *            back each ancestor is pruned if its 8 children are equal-valued leaves, else set to the max child
 *   .bt      writeBinary = toMaxLikelihood + prune + 2 bytes per inner node (2 bits per child:
 *            bit0/bit1 = 1/0 free leaf, 0/1 occupied leaf, 1/1 inner, 0/0 unknown)
 * PARITY UNPINNED by the reference for keys/log-odds/prune (octomap_mesh_masking_unit.cpp:59-70 checks non-null
 * only); the .bt codec is pinned by a byte-exact round trip of the reference fixture box_2m.bt.
 *
 * Derivation, rule by rule (upstream function -> the expression restated here; upstream = octomap 1.9.x / tesseract
 * 0.1-era headers as remembered, neither is on this machine, so none of this can be re-read offline):
 *   resolution_factor   OcTreeBaseImpl::setResolution:      resolution_factor = 1. / resolution            (double)
 *   key                 OcTreeBaseImpl::coordToKeyChecked:  int scaled = ((int) floor(resolution_factor * coordinate)) +
 *                       tree_max_val; valid iff scaled >= 0 && (unsigned) scaled < 2 * tree_max_val; the float coordinate of a
 *                       pcl point is promoted to double first (updateNode(double x, double y, double z, bool))
 *                       -> orc_coord_to_key: floor(inv_res * (double)x) + 32768, product rounded once in double
 *                       (tests/test_oracle_keys.py enumerates every voxel boundary of three resolutions against exact
 *                       rational arithmetic, for this file and, in test_gpu_keys.py, for the CUDA path)
 *   hit / clamp values  octomap_utils.h logodds(double p) = (float) log(p / (1 - p)): DOUBLE log, then one cast to float;
 *                       AbstractOccupancyOcTree(): setProbHit(0.7), setClampingThresMax(0.971), setClampingThresMin(0.1192)
 *                       -> 0x3f58e883 and 0x4060b4ba; evaluating in float (logf) would give 0x3f58e882 / 0x4060b4bc,
 *                       which tests/test_oracle_keys.py shows is NOT what the LUT of the CUDA path holds
 *   leaf update         OccupancyOcTreeBase::updateNodeLogOdds: node->addValue(update) (float +=); if (< clamping_thres_min)
 *                       set min and return; if (> clamping_thres_max) set max  -> hits only: 0x3f58e883, 0x3fd8e883,
 *                       0x4022ae62, 0x4058e883, then 0x4060b4ba for ever
 *   early return        OccupancyOcTreeBase::updateNode(key, float log_odds_update, lazy): leaf = search(key); if (leaf &&
 *                       log_odds_update >= 0 && leaf->getLogOdds() >= clamping_thres_max) return leaf
 *   inner nodes         updateNodeRecurs: after the child returns, if (pruneNode(node)) ... else node->updateOccupancyChildren()
 *                       (= max of the children's log-odds); pruneNode needs 8 existing, childless, equal-valued children
 *   tesseract           tesseract_geometry::Octree(cloud, resolution, sub_type, prune): new octomap::OcTree(resolution); for each
 *                       point updateNode(x, y, z, true); then, if prune, tesseract's own prune(): an inner node whose 8
 *                       children are leaves ABOVE the occupancy threshold collapses (values need not be equal, the parent
 *                       keeps child 0's value) -> orc_octree_finalize_tesseract
 * What stays unpinned after the tests above: that these remembered expressions ARE upstream's (in particular tesseract's
 * prune rule and the early return), and the operation order of pcl::transformPointCloud (oracle/cloud.c).
 * This tree is also the timed CPU baseline for integration (16-level pointer descent, `malloc` per node).
 */
#include "pcs_oracle.h"

#include <float.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define TREE_DEPTH 16
#define TREE_MAX_VAL 32768

typedef struct node
{
  float value;
  struct node** children; /* NULL or array of 8 */
} node;

struct orc_octree
{
  node* root;
  double resolution;
  double resolution_factor;
  float prob_hit_log, clamp_min, clamp_max, occ_thres;
  size_t tree_size;
};

static float logodds(double p) { return (float)log(p / (1.0 - p)); }

orc_octree* orc_octree_create(double resolution)
{
  orc_octree* t = (orc_octree*)calloc(1, sizeof(orc_octree));
  t->resolution = resolution;
  t->resolution_factor = 1.0 / resolution;
  t->prob_hit_log = logodds(0.7);
  t->clamp_min = logodds(0.1192);
  t->clamp_max = logodds(0.971);
  t->occ_thres = logodds(0.5);
  return t;
}

double orc_octree_resolution(const orc_octree* t) { return t->resolution; }

static void free_node(node* n)
{
  if (!n)
    return;
  if (n->children)
  {
    for (int i = 0; i < 8; i++)
      free_node(n->children[i]);
    free(n->children);
  }
  free(n);
}

void orc_octree_destroy(orc_octree* t)
{
  if (!t)
    return;
  free_node(t->root);
  free(t);
}

static node* new_node(orc_octree* t)
{
  node* n = (node*)malloc(sizeof(node));
  n->value = 0.0f;
  n->children = NULL;
  t->tree_size++;
  return n;
}

static int child_exists(const node* n, int i) { return n->children && n->children[i]; }

static int has_children(const node* n)
{
  if (!n->children)
    return 0;
  for (int i = 0; i < 8; i++)
    if (n->children[i])
      return 1;
  return 0;
}

static node* create_child(orc_octree* t, node* n, int i)
{
  if (!n->children)
    n->children = (node**)calloc(8, sizeof(node*));
  n->children[i] = new_node(t);
  return n->children[i];
}

static void expand_node(orc_octree* t, node* n)
{
  for (int i = 0; i < 8; i++)
    create_child(t, n, i)->value = n->value;
}

static void delete_children(orc_octree* t, node* n)
{
  for (int i = 0; i < 8; i++)
  {
    free(n->children[i]);
    t->tree_size--;
  }
  free(n->children);
  n->children = NULL;
}

static int collapsible_equal(const node* n)
{
  if (!child_exists(n, 0))
    return 0;
  const node* first = n->children[0];
  if (has_children(first))
    return 0;
  for (int i = 1; i < 8; i++)
    if (!child_exists(n, i) || has_children(n->children[i]) || !(n->children[i]->value == first->value))
      return 0;
  return 1;
}

static int prune_node_equal(orc_octree* t, node* n)
{
  if (!collapsible_equal(n))
    return 0;
  n->value = n->children[0]->value;
  delete_children(t, n);
  return 1;
}

static float max_child(const node* n)
{
  float m = -FLT_MAX;
  if (n->children)
    for (int i = 0; i < 8; i++)
      if (n->children[i] && n->children[i]->value > m)
        m = n->children[i]->value;
  return m;
}

static inline int child_idx(const uint16_t k[3], int level)
{
  int pos = 0;
  if (k[0] & (1 << level))
    pos += 1;
  if (k[1] & (1 << level))
    pos += 2;
  if (k[2] & (1 << level))
    pos += 4;
  return pos;
}

static const node* search_key(const orc_octree* t, const uint16_t k[3], int* depth_out)
{
  if (!t->root)
    return NULL;
  const node* cur = t->root;
  for (int level = TREE_DEPTH - 1; level >= 0; level--)
  {
    const int pos = child_idx(k, level);
    if (child_exists(cur, pos))
      cur = cur->children[pos];
    else
    {
      if (!has_children(cur))
      {
        if (depth_out)
          *depth_out = TREE_DEPTH - 1 - level;
        return cur; /* pruned leaf at a coarser level */
      }
      return NULL;
    }
  }
  if (depth_out)
    *depth_out = TREE_DEPTH;
  return cur;
}

static node* update_recurs(orc_octree* t, node* n, int just_created, const uint16_t k[3], int depth, float upd)
{
  int created = 0;
  if (depth < TREE_DEPTH)
  {
    const int pos = child_idx(k, TREE_DEPTH - 1 - depth);
    if (!child_exists(n, pos))
    {
      if (!has_children(n) && !just_created)
        expand_node(t, n);
      else
      {
        create_child(t, n, pos);
        created = 1;
      }
    }
    node* ret = update_recurs(t, n->children[pos], created, k, depth + 1, upd);
    if (prune_node_equal(t, n))
      ret = n;
    else
      n->value = max_child(n);
    return ret;
  }
  n->value += upd;
  if (n->value < t->clamp_min)
    n->value = t->clamp_min;
  if (n->value > t->clamp_max)
    n->value = t->clamp_max;
  return n;
}

void orc_octree_insert_key(orc_octree* t, uint16_t kx, uint16_t ky, uint16_t kz)
{
  const uint16_t k[3] = { kx, ky, kz };
  const float upd = t->prob_hit_log;
  const node* leaf = search_key(t, k, NULL);
  if (leaf && upd >= 0 && leaf->value >= t->clamp_max)
    return; /* early abort, no change will happen */
  int created_root = 0;
  if (!t->root)
  {
    t->root = new_node(t);
    created_root = 1;
  }
  update_recurs(t, t->root, created_root, k, 0, upd);
}

int orc_coord_to_key(double inv_res, double coord, uint16_t* key)
{
  /* octomap relies on (int) of an out-of-range / NaN double (INT_MIN on x86) failing the range test; stated
   * explicitly here instead of through undefined behaviour (SURVEY 7 "NaN / range parity"). */
  const double s = floor(inv_res * coord);
  if (!(s >= -(double)TREE_MAX_VAL && s < (double)TREE_MAX_VAL))
    return 0;
  *key = (uint16_t)((int)s + TREE_MAX_VAL);
  return 1;
}

int orc_octree_insert_point(orc_octree* t, float x, float y, float z)
{
  uint16_t k[3];
  if (!orc_coord_to_key(t->resolution_factor, (double)x, &k[0]) ||
      !orc_coord_to_key(t->resolution_factor, (double)y, &k[1]) ||
      !orc_coord_to_key(t->resolution_factor, (double)z, &k[2]))
    return 0;
  orc_octree_insert_key(t, k[0], k[1], k[2]);
  return 1;
}

void orc_octree_insert_cloud(orc_octree* t, const orc_point* pts, size_t n)
{
  for (size_t i = 0; i < n; i++)
    orc_octree_insert_point(t, pts[i].x, pts[i].y, pts[i].z);
}

static void inner_occ_recurs(node* n, int depth)
{
  if (has_children(n))
  {
    if (depth < TREE_DEPTH)
      for (int i = 0; i < 8; i++)
        if (child_exists(n, i))
          inner_occ_recurs(n->children[i], depth + 1);
    n->value = max_child(n);
  }
}

void orc_octree_update_inner_occupancy(orc_octree* t)
{
  if (t->root)
    inner_occ_recurs(t->root, 0);
}

/* tesseract_geometry::Octree::prune: all 8 children exist, are leaves and are occupied */
static int collapsible_occupied(const orc_octree* t, const node* n)
{
  if (!child_exists(n, 0))
    return 0;
  for (int i = 0; i < 8; i++)
    if (!child_exists(n, i) || has_children(n->children[i]) || n->children[i]->value < t->occ_thres)
      return 0;
  return 1;
}

static void prune_recurs(orc_octree* t, node* n, int depth, int max_depth, unsigned* num_pruned, int tesseract)
{
  if (depth < max_depth)
  {
    for (int i = 0; i < 8; i++)
      if (child_exists(n, i))
        prune_recurs(t, n->children[i], depth + 1, max_depth, num_pruned, tesseract);
  }
  else if (tesseract)
  {
    if (collapsible_occupied(t, n))
    {
      n->value = n->children[0]->value;
      delete_children(t, n);
      (*num_pruned)++;
    }
  }
  else if (prune_node_equal(t, n))
    (*num_pruned)++;
}

static void prune_impl(orc_octree* t, int tesseract)
{
  if (!t->root)
    return;
  for (int depth = TREE_DEPTH - 1; depth > 0; --depth)
  {
    unsigned num_pruned = 0;
    prune_recurs(t, t->root, 0, depth, &num_pruned, tesseract);
    if (num_pruned == 0)
      break;
  }
}

void orc_octree_prune_tesseract(orc_octree* t) { prune_impl(t, 1); }
void orc_octree_prune_octomap(orc_octree* t) { prune_impl(t, 0); }

static void ml_recurs(orc_octree* t, node* n, int depth, int max_depth)
{
  if (depth < max_depth)
  {
    for (int i = 0; i < 8; i++)
      if (child_exists(n, i))
        ml_recurs(t, n->children[i], depth + 1, max_depth);
  }
  else
    n->value = (n->value >= t->occ_thres) ? t->clamp_max : t->clamp_min;
}

void orc_octree_to_max_likelihood(orc_octree* t)
{
  if (!t->root)
    return;
  for (int depth = TREE_DEPTH; depth > 0; depth--)
    ml_recurs(t, t->root, 0, depth);
  t->root->value = (t->root->value >= t->occ_thres) ? t->clamp_max : t->clamp_min;
}

size_t orc_octree_num_nodes(const orc_octree* t) { return t->tree_size; }

typedef struct
{
  uint16_t *kx, *ky, *kz;
  uint8_t* depth;
  float* lo;
  size_t cap, n;
} leaf_sink;

static void leaves_recurs(const node* n, int depth, uint32_t x, uint32_t y, uint32_t z, leaf_sink* s)
{
  if (!has_children(n))
  {
    if (s->n < s->cap)
    {
      if (s->kx)
      {
        s->kx[s->n] = (uint16_t)x;
        s->ky[s->n] = (uint16_t)y;
        s->kz[s->n] = (uint16_t)z;
      }
      if (s->depth)
        s->depth[s->n] = (uint8_t)depth;
      if (s->lo)
        s->lo[s->n] = n->value;
    }
    s->n++;
    return;
  }
  const uint32_t half = 1u << (TREE_DEPTH - 1 - depth);
  for (int i = 0; i < 8; i++)
    if (n->children[i])
      leaves_recurs(n->children[i], depth + 1, x + ((i & 1) ? half : 0), y + ((i & 2) ? half : 0),
                    z + ((i & 4) ? half : 0), s);
}

size_t orc_octree_leaves(const orc_octree* t, uint16_t* kx, uint16_t* ky, uint16_t* kz, uint8_t* depth, float* logodds_out,
                         size_t cap)
{
  leaf_sink s = { kx, ky, kz, depth, logodds_out, cap, 0 };
  if (t->root)
    leaves_recurs(t->root, 0, 0, 0, 0, &s);
  return s.n;
}

size_t orc_octree_num_leaves(const orc_octree* t) { return orc_octree_leaves(t, NULL, NULL, NULL, NULL, NULL, 0); }

int orc_octree_search(const orc_octree* t, uint16_t kx, uint16_t ky, uint16_t kz, float* lo, int* depth)
{
  const uint16_t k[3] = { kx, ky, kz };
  int d = 0;
  const node* n = search_key(t, k, &d);
  if (!n)
    return 0;
  if (lo)
    *lo = n->value;
  if (depth)
    *depth = d;
  return 1;
}

float orc_logodds_after_hits(uint32_t hits)
{
  const float hit = logodds(0.7), mx = logodds(0.971);
  float v = 0.0f;
  for (uint32_t i = 0; i < hits; i++)
  {
    if (v >= mx)
      break;
    v += hit;
    if (v > mx)
      v = mx;
  }
  return v;
}

/* ---- .bt codec ----------------------------------------------------------------------------------- */
typedef struct
{
  const uint8_t* p;
  size_t n, pos;
  int fail;
} rd;

static void read_node(orc_octree* t, node* n, rd* r)
{
  if (r->pos + 2 > r->n)
  {
    r->fail = 1;
    return;
  }
  const unsigned bits = r->p[r->pos] | ((unsigned)r->p[r->pos + 1] << 8);
  r->pos += 2;
  n->value = t->clamp_max; /* inner nodes default to occupied */
  for (int i = 0; i < 8; i++)
  {
    const unsigned c = (bits >> (2 * i)) & 3u; /* bit0 = child[i*2], bit1 = child[i*2+1] */
    if (c == 1u)
      create_child(t, n, i)->value = t->clamp_min; /* free leaf */
    else if (c == 2u)
      create_child(t, n, i)->value = t->clamp_max; /* occupied leaf */
    else if (c == 3u)
      create_child(t, n, i)->value = -200.0f; /* has children */
  }
  for (int i = 0; i < 8; i++)
    if (child_exists(n, i) && fabs(n->children[i]->value + 200.0) < 1e-3)
    {
      read_node(t, n->children[i], r);
      if (r->fail)
        return;
      n->children[i]->value = max_child(n->children[i]);
    }
}

orc_octree* orc_octree_read_bt(const uint8_t* bytes, size_t n)
{
  /* header: comment lines, then "id OcTree", "size N", "res R", "data" */
  size_t pos = 0;
  double res = 0.0;
  int have_data = 0;
  while (pos < n && !have_data)
  {
    size_t e = pos;
    while (e < n && bytes[e] != '\n')
      e++;
    char line[256];
    size_t len = e - pos < sizeof(line) - 1 ? e - pos : sizeof(line) - 1;
    memcpy(line, bytes + pos, len);
    line[len] = 0;
    pos = e + 1;
    if (line[0] == '#')
      continue;
    if (!strncmp(line, "res ", 4))
      res = atof(line + 4);
    else if (!strncmp(line, "data", 4))
      have_data = 1;
  }
  if (!have_data || res <= 0.0)
    return NULL;
  orc_octree* t = orc_octree_create(res);
  if (pos < n)
  {
    t->root = new_node(t);
    rd r = { bytes, n, pos, 0 };
    read_node(t, t->root, &r);
    if (r.fail)
    {
      orc_octree_destroy(t);
      return NULL;
    }
  }
  return t;
}

typedef struct
{
  uint8_t* p;
  size_t cap, n;
} wr;

static void put(wr* w, uint8_t b)
{
  if (w->n < w->cap)
    w->p[w->n] = b;
  w->n++;
}

static void write_node(const orc_octree* t, const node* n, wr* w)
{
  unsigned bits = 0;
  for (int i = 0; i < 8; i++)
    if (child_exists(n, i))
    {
      const node* c = n->children[i];
      unsigned v;
      if (has_children(c))
        v = 3u;
      else if (c->value >= t->occ_thres)
        v = 2u;
      else
        v = 1u;
      bits |= v << (2 * i);
    }
  put(w, (uint8_t)(bits & 0xff));
  put(w, (uint8_t)(bits >> 8));
  for (int i = 0; i < 8; i++)
    if (child_exists(n, i) && has_children(n->children[i]))
      write_node(t, n->children[i], w);
}

size_t orc_octree_write_bt(orc_octree* t, uint8_t* bytes, size_t cap)
{
  orc_octree_to_max_likelihood(t);
  orc_octree_prune_octomap(t);
  char hdr[512];
  const int hl = snprintf(hdr, sizeof(hdr),
                          "# Octomap OcTree binary file\n# (feel free to add / change comments, but leave the first "
                          "line as it is!)\n#\nid OcTree\nsize %zu\nres %g\ndata\n",
                          t->tree_size, t->resolution);
  wr w = { bytes, cap, 0 };
  for (int i = 0; i < hl; i++)
    put(&w, (uint8_t)hdr[i]);
  if (t->root)
    write_node(t, t->root, &w);
  return w.n;
}

/* ---- flat reference map ---------------------------------------------------------------------------- */
uint64_t orc_morton48(uint16_t kx, uint16_t ky, uint16_t kz)
{
  uint64_t m = 0;
  for (int i = 0; i < 16; i++)
  {
    m |= (uint64_t)((kx >> i) & 1u) << (3 * i);
    m |= (uint64_t)((ky >> i) & 1u) << (3 * i + 1);
    m |= (uint64_t)((kz >> i) & 1u) << (3 * i + 2);
  }
  return m;
}

typedef struct
{
  uint64_t morton, key;
  uint32_t cls;
} krec;

static int krec_cmp(const void* a, const void* b)
{
  const krec* x = (const krec*)a;
  const krec* y = (const krec*)b;
  if (x->morton != y->morton)
    return x->morton < y->morton ? -1 : 1;
  return 0;
}

size_t orc_count_voxels(const orc_point* pts, size_t n, const float* pose, double resolution, int lower, int upper,
                        int limit_negative, uint64_t* key48, uint32_t* count0, uint32_t* count1, size_t cap)
{
  const double inv = 1.0 / resolution;
  krec* recs = (krec*)malloc(sizeof(krec) * (n ? n : 1));
  size_t m = 0;
  for (size_t i = 0; i < n; i++)
  {
    orc_point p = pts[i];
    if (pose)
      orc_transform_points(pose, &p, 1, &p);
    uint16_t k[3];
    if (!orc_coord_to_key(inv, (double)p.x, &k[0]) || !orc_coord_to_key(inv, (double)p.y, &k[1]) ||
        !orc_coord_to_key(inv, (double)p.z, &k[2]))
      continue;
    recs[m].morton = orc_morton48(k[0], k[1], k[2]);
    recs[m].key = ((uint64_t)k[2] << 32) | ((uint64_t)k[1] << 16) | k[0];
    recs[m].cls = (uint32_t)orc_color_pass(p.r, p.g, p.b, lower, upper, limit_negative);
    m++;
  }
  qsort(recs, m, sizeof(krec), krec_cmp);
  size_t nk = 0;
  for (size_t i = 0; i < m;)
  {
    size_t j = i;
    uint32_t c0 = 0, c1 = 0;
    while (j < m && recs[j].morton == recs[i].morton)
    {
      if (recs[j].cls)
        c1++;
      else
        c0++;
      j++;
    }
    if (nk < cap)
    {
      key48[nk] = recs[i].key;
      count0[nk] = c0;
      count1[nk] = c1;
    }
    nk++;
    i = j;
  }
  free(recs);
  return nk;
}
