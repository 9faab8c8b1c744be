/*
 * pcs_oracle.h -- CPU restatement ("oracle") of the annotate-and-integrate hot path of
 * swri-robotics/point_cloud_segmentation.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product: only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it,
 * and there only as the checker or as the timed CPU baseline.  The product path
 * (point_cloud_segmentation_b200/csrc) never links, imports or falls back to this code.
 *
 * The reference's own C++ cannot be built in this environment (OpenCV C++, PCL, octomap,
 * tesseract, Bullet, Eigen, assimp, console_bridge and gtest are all absent), so every
 * function here restates the reference file:line it follows, and the third-party
 * arithmetic behind it (OpenCV 3 cvtColor/inRange, octomap OcTree::updateNode,
 * tesseract_geometry::Octree's prune, Bullet box/triangle contact at distance <= 0).
 *
 * Pinning status (see tests/test_oracle_*.py):
 *   HSV + inRange          pinned against cv2 4.13 on all 2^24 colours (reference leaves it unpinned)
 *   cloudToImage/imageToCloud  pinned by the reference's round-trip test (utils_unit.cpp:38-64)
 *   colorPassthrough       pinned by octomap_mesh_masking_unit.cpp:90-135 (exact counts)
 *   mesh mask              pinned by box_2m.bt x plane_4m.stl: OUTSIDE 11400, COLORIZED 12288/6273;
 *                          INSIDE is a Bullet contact multiset (1975) -- geometry gives 888 unique
 *                          triangles / 1976 (pruned leaf, triangle) pairs; set semantics are pinned,
 *                          Bullet's duplicate count is not reproducible.
 *   .bt reader/writer      pinned by byte-exact round trip of box_2m.bt
 *   octree-from-cloud (keys, log-odds, prune) and pcl transform: PARITY UNPINNED by the
 *                          reference; restated from octomap / tesseract / PCL upstream rules.
 */
#ifndef PCS_ORACLE_H
#define PCS_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* pcl::PointXYZRGB memory layout (32 B, 16-B aligned): x,y,z, data[3], then b,g,r,a, then padding. */
typedef struct orc_point
{
  float x, y, z, w;
  uint8_t b, g, r, a;
  uint32_t pad[3];
} orc_point;

typedef struct orc_hsv_params
{
  int32_t h_lo, h_hi, s_lo, s_hi, v_lo, v_hi; /* inclusive bounds, hsv_thresholding.h:47-52 */
  int32_t inverted;                           /* hsv_thresholding.h:61-62 */
  int32_t morph_open;                         /* 0 = off (reference behaviour); 1 = 5x5 MORPH_OPEN */
} orc_hsv_params;

void orc_hsv_default_params(orc_hsv_params* p);

/* ---- stage 1: detector ---------------------------------------------------------------- */
void orc_hsv_tables(int32_t sdiv[256], int32_t hdiv[256]);
void orc_bgr2hsv(const uint8_t* bgr, int w, int h, size_t stride, uint8_t* hsv, size_t hsv_stride);
void orc_hsv_mask(const uint8_t* bgr, int w, int h, size_t stride, const orc_hsv_params* p, uint8_t* mask,
                  size_t mask_stride);
void orc_morph_open5(const uint8_t* in, int w, int h, uint8_t* out);
/* hash of the mask over all 2^24 BGR colours (b fastest), FNV-1a 64 */
uint64_t orc_hsv_exhaustive_hash(const orc_hsv_params* p, uint64_t* n_positive);

/* histogram back-projection detector (pcs_detection/src_python/pcs_detection/models/histogram_backprojection.py:22-52):
 * BGR2HSV, calcBackProject on (H,S) with ranges [0,180),[0,255), 5x5 ellipse filter2D, threshold, 5x5 open, min-max
 * normalise to {0,1}, x255, 3 channels.  hist is h_bins x s_bins float32, row-major. */
void orc_backproject_tables(int h_bins, int s_bins, int32_t tab_h[256], int32_t tab_s[256]);
void orc_histogram_backprojection(const uint8_t* bgr, int w, int h, size_t stride, const float* hist, int h_bins,
                                  int s_bins, int thr_min, int thr_value, uint8_t* out_bgr3);

/* ---- stage 2: cloud <-> image, labelling ---------------------------------------------- */
void orc_cloud_to_image(const orc_point* cloud, int w, int h, double* position_image, uint8_t* bgr_image);
void orc_image_to_cloud(const uint8_t* annotation_bgr, const double* position_image, int w, int h, orc_point* out);
size_t orc_color_passthrough(const orc_point* in, size_t n, int lower, int upper, int limit_negative,
                             orc_point* out);
int orc_color_pass(uint8_t r, uint8_t g, uint8_t b, int lower, int upper, int limit_negative);
/* float 4x4 row-major T; ((m0*x + m1*y) + m2*z) + m3 with no FMA contraction */
void orc_transform_points(const float* T, const orc_point* in, size_t n, orc_point* out);
/* Full reference-shaped annotate of one organized frame:
 * cloudToImage -> hsvThresholdingDetector -> merge(mask x3) -> imageToCloud. */
void orc_annotate_frame(const orc_point* cloud, int w, int h, const orc_hsv_params* p, orc_point* annotated,
                        uint8_t* mask);

/* ---- stage 3: octomap-style pointer octree -------------------------------------------- */
typedef struct orc_octree orc_octree;
orc_octree* orc_octree_create(double resolution);
void orc_octree_destroy(orc_octree* t);
double orc_octree_resolution(const orc_octree* t);
/* octomap coordToKeyChecked; returns 0 if out of range / non-finite */
int orc_coord_to_key(double inv_res, double coord, uint16_t* key);
/* updateNode(x,y,z,true); returns 1 if the point produced a key (was not skipped) */
int orc_octree_insert_point(orc_octree* t, float x, float y, float z);
void orc_octree_insert_key(orc_octree* t, uint16_t kx, uint16_t ky, uint16_t kz);
void orc_octree_insert_cloud(orc_octree* t, const orc_point* pts, size_t n);
void orc_octree_update_inner_occupancy(orc_octree* t);
void orc_octree_prune_tesseract(orc_octree* t); /* collapse complete occupied octets, break-if-none quirk */
void orc_octree_prune_octomap(orc_octree* t);   /* octomap::prune(): equal-valued octets */
void orc_octree_to_max_likelihood(orc_octree* t);
size_t orc_octree_num_nodes(const orc_octree* t);
size_t orc_octree_num_leaves(const orc_octree* t);
/* leaves in child-index DFS order; depth 16 = finest. keys are the min-corner key of the leaf. */
size_t orc_octree_leaves(const orc_octree* t, uint16_t* kx, uint16_t* ky, uint16_t* kz, uint8_t* depth,
                         float* logodds, size_t cap);
/* search(); returns 1 and the leaf's log-odds / depth if found */
int orc_octree_search(const orc_octree* t, uint16_t kx, uint16_t ky, uint16_t kz, float* logodds, int* depth);
/* .bt (octomap binary) */
orc_octree* orc_octree_read_bt(const uint8_t* bytes, size_t n);
size_t orc_octree_write_bt(orc_octree* t, uint8_t* bytes, size_t cap); /* does toMaxLikelihood+prune first */
float orc_logodds_after_hits(uint32_t hits);

/* flat reference map: exact per-voxel counts, sorted by Morton order.  key48 = kz<<32 | ky<<16 | kx. */
uint64_t orc_morton48(uint16_t kx, uint16_t ky, uint16_t kz);
size_t orc_count_voxels(const orc_point* pts, size_t n, const float* pose, double resolution, int lower, int upper,
                        int limit_negative, uint64_t* key48, uint32_t* count0, uint32_t* count1, size_t cap);

/* ---- stage 4: mesh mask --------------------------------------------------------------- */
/* closed (touching counts) triangle / axis-aligned box overlap on doubles, 13-axis SAT */
int orc_tri_box_overlap(const double lo[3], const double hi[3], const double v0[3], const double v1[3],
                        const double v2[3]);
/* flags[t] = number of (leaf box, triangle) overlapping pairs for triangle t (0 = outside).
 * Boxes are the occupied (logodds >= 0) leaves of the tree at their own depth (pruned leaves are big boxes). */
uint64_t orc_mesh_contacts_tree(const orc_octree* t, const double* verts, size_t nv, const int32_t* faces4,
                                size_t nt, uint32_t* pair_count);
/* same, from a flat occupied key list at leaf resolution (set semantics; pair_count is per finest voxel) */
uint64_t orc_mesh_contacts_keys(const uint64_t* key48_sorted_packed, size_t nk, double resolution,
                                const double* verts, size_t nv, const int32_t* faces4, size_t nt,
                                uint8_t* tri_flags);
/* brute force over all (box,triangle) pairs, for validating the range logic on small inputs */
uint64_t orc_mesh_contacts_brute(const double* box_lo, const double* box_hi, size_t nb, const double* verts,
                                 const int32_t* faces4, size_t nt, uint32_t* pair_count);
/* maskMesh assembly, octomap_mesh_masking.cpp:163-280. mask_type 0 INSIDE(unique, input order) 1 OUTSIDE 2 COLORIZED */
size_t orc_mask_mesh_emit(const double* verts, size_t nv, const int32_t* faces4, size_t nt,
                          const uint8_t* tri_flags, int mask_type, double* out_verts, int32_t* out_faces4,
                          uint8_t* out_rgb, size_t* n_out_verts);

/* mesh I/O (assimp-equivalent: triangulated, identical vertices welded) */
int orc_load_stl_binary(const uint8_t* bytes, size_t n, double* verts, size_t vcap, int32_t* faces4, size_t tcap,
                        size_t* nv, size_t* nt);
int orc_load_ply_ascii(const char* text, size_t n, double* verts, size_t vcap, int32_t* faces4, size_t tcap,
                       size_t* nv, size_t* nt);
size_t orc_write_ply_ascii(const double* verts, size_t nv, const uint8_t* rgb, const int32_t* faces4, size_t nt,
                           char* out, size_t cap);

/* ---- synthetic scan -------------------------------------------------------- */
uint64_t orc_splitmix64(uint64_t x);
double orc_psin(double x);
double orc_pcos(double x);
void orc_synth_pose(int frame, float pose[16]);
void orc_synth_frame(uint64_t seed, int frame, int w, int h, const float pose[16], orc_point* out);

/* ---- timed CPU baseline: the whole reference-shaped path over a batch of frames --------- */
typedef struct orc_baseline_result
{
  double seconds_annotate;
  double seconds_integrate;
  uint64_t points;
  uint64_t inserted;
  uint64_t leaves;
} orc_baseline_result;
/* threads > 1 shards the annotate stage by frame (pthreads); integration stays serial like octomap. */
void orc_baseline_annotate_integrate(const orc_point* frames, int n_frames, int w, int h, const float* poses,
                                     const orc_hsv_params* p, double resolution, int lower, int upper,
                                     int limit_negative, int threads, orc_baseline_result* res);

#ifdef __cplusplus
}
#endif
#endif
