/*
 * oracle/cloud.c -- TEST INFRASTRUCTURE (see pcs_oracle.h).  CPU restatement of the cloud <-> image utilities,
 * the label filter and the rigid transform on the annotate-and-integrate path.
 *
 *   orc_cloud_to_image     pcs_detection/include/pcs_detection/utils.h:98-129
 *   orc_image_to_cloud     pcs_detection/include/pcs_detection/utils.h:138-171
 *   orc_annotate_frame     point_cloud_annotator.cpp:7-82 (one frame) with the detector service of
 *                          pcs_ros/src/hsv_thresholding_node.cpp:8-28 (mask replicated to 3 channels, :20-23)
 *   orc_color_passthrough  pcs_scan_integration/src/pcs_scan_integration/octomap_mesh_masking.cpp:52-96
 *   orc_transform_points   pcs_ros/src/octomap_mesh_mask_node.cpp:52-57 (pcl_ros::transformPointCloud; PCL,
 *                          not in the tree: float row-by-point products summed left to right, non-finite
 *                          points copied through unchanged).  PARITY UNPINNED by the reference.
 * Build with -ffp-contract=off so no product/sum here is fused.
 */
#include "pcs_oracle.h"

#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

void orc_cloud_to_image(const orc_point* cloud, int w, int h, double* position_image, uint8_t* bgr_image)
{
  for (int y = 0; y < h; y++)
    for (int x = 0; x < w; x++)
    {
      const orc_point* p = &cloud[(size_t)y * w + x];
      double* pos = position_image + ((size_t)y * w + x) * 3;
      pos[0] = p->x; /* utils.h:117-119, float -> double widen (exact) */
      pos[1] = p->y;
      pos[2] = p->z;
      uint8_t* c = bgr_image + ((size_t)y * w + x) * 3;
      c[0] = p->b; /* utils.h:122-124: B,G,R order */
      c[1] = p->g;
      c[2] = p->r;
    }
}

void orc_image_to_cloud(const uint8_t* annotation_bgr, const double* position_image, int w, int h, orc_point* out)
{
  for (int y = 0; y < h; y++)
    for (int x = 0; x < w; x++)
    {
      const size_t i = (size_t)y * w + x;
      orc_point pt;
      /* pcl::PointXYZRGB(): x=y=z=0, data[3]=1.f, r=g=b=0, a=255; padding made deterministic (zero) here */
      memset(&pt, 0, sizeof(pt));
      pt.w = 1.0f;
      pt.x = (float)position_image[3 * i + 0]; /* utils.h:152-154 */
      pt.y = (float)position_image[3 * i + 1];
      pt.z = (float)position_image[3 * i + 2];
      const uint8_t r = annotation_bgr[3 * i + 2], g = annotation_bgr[3 * i + 1], b = annotation_bgr[3 * i + 0];
      const int32_t rgb = (r << 16) | (g << 8) | b; /* utils.h:161: alpha byte becomes 0 */
      memcpy(&pt.b, &rgb, 4);
      out[i] = pt;
    }
}

int orc_color_pass(uint8_t r, uint8_t g, uint8_t b, int lower, int upper, int limit_negative)
{
  if (!limit_negative) /* octomap_mesh_masking.cpp:75-77 */
    return (r >= lower && r <= upper) && (g >= lower && g <= upper) && (b >= lower && b <= upper);
  /* :85-87 */
  return (r <= lower || r >= upper) && (g <= lower || g >= upper) && (b <= lower || b >= upper);
}

size_t orc_color_passthrough(const orc_point* in, size_t n, int lower, int upper, int limit_negative, orc_point* out)
{
  size_t k = 0;
  for (size_t i = 0; i < n; i++)
    if (orc_color_pass(in[i].r, in[i].g, in[i].b, lower, upper, limit_negative))
      out[k++] = in[i];
  return k;
}

void orc_transform_points(const float* T, const orc_point* in, size_t n, orc_point* out)
{
  for (size_t i = 0; i < n; i++)
  {
    orc_point p = in[i];
    if (isfinite(p.x) && isfinite(p.y) && isfinite(p.z))
    {
      const float x = p.x, y = p.y, z = p.z;
      p.x = ((T[0] * x + T[1] * y) + T[2] * z) + T[3];
      p.y = ((T[4] * x + T[5] * y) + T[6] * z) + T[7];
      p.z = ((T[8] * x + T[9] * y) + T[10] * z) + T[11];
    }
    out[i] = p;
  }
}

void orc_annotate_frame(const orc_point* cloud, int w, int h, const orc_hsv_params* p, orc_point* annotated,
                        uint8_t* mask)
{
  const size_t n = (size_t)w * h;
  double* pos = (double*)malloc(n * 3 * sizeof(double));
  uint8_t* bgr = (uint8_t*)malloc(n * 3);
  uint8_t* m = mask ? mask : (uint8_t*)malloc(n);
  uint8_t* ann = (uint8_t*)malloc(n * 3);
  orc_cloud_to_image(cloud, w, h, pos, bgr);         /* point_cloud_annotator.cpp:12 */
  orc_hsv_mask(bgr, w, h, (size_t)w * 3, p, m, (size_t)w); /* hsv_thresholding_node.cpp:17 */
  for (size_t i = 0; i < n; i++)                      /* cv::merge, hsv_thresholding_node.cpp:20-22 */
    ann[3 * i] = ann[3 * i + 1] = ann[3 * i + 2] = m[i];
  orc_image_to_cloud(ann, pos, w, h, annotated); /* point_cloud_annotator.cpp:65-66 */
  free(ann);
  if (!mask)
    free(m);
  free(bgr);
  free(pos);
}

/* ---- timed baseline ------------------------------------------------------------------------------ */
static double now_s(void)
{
  struct timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

typedef struct
{
  const orc_point* frames;
  orc_point* annotated;
  int w, h, n_frames, tid, threads;
  const float* poses;
  const orc_hsv_params* p;
} ann_job;

static void* ann_worker(void* arg)
{
  ann_job* j = (ann_job*)arg;
  const size_t n = (size_t)j->w * j->h;
  for (int f = j->tid; f < j->n_frames; f += j->threads)
  {
    orc_point* out = j->annotated + (size_t)f * n;
    orc_annotate_frame(j->frames + (size_t)f * n, j->w, j->h, j->p, out, NULL);
    if (j->poses)
      orc_transform_points(j->poses + (size_t)f * 16, out, n, out);
  }
  return NULL;
}

void orc_baseline_annotate_integrate(const orc_point* frames, int n_frames, int w, int h, const float* poses,
                                     const orc_hsv_params* p, double resolution, int lower, int upper,
                                     int limit_negative, int threads, orc_baseline_result* res)
{
  const size_t n = (size_t)w * h;
  if (threads < 1)
    threads = 1;
  orc_point* annotated = (orc_point*)malloc(n * (size_t)n_frames * sizeof(orc_point));
  memset(res, 0, sizeof(*res));

  double t0 = now_s();
  if (threads == 1)
  {
    ann_job j = { frames, annotated, w, h, n_frames, 0, 1, poses, p };
    ann_worker(&j);
  }
  else
  {
    pthread_t* th = (pthread_t*)malloc(sizeof(pthread_t) * (size_t)threads);
    ann_job* jobs = (ann_job*)malloc(sizeof(ann_job) * (size_t)threads);
    for (int t = 0; t < threads; t++)
    {
      ann_job j = { frames, annotated, w, h, n_frames, t, threads, poses, p };
      jobs[t] = j;
      pthread_create(&th[t], NULL, ann_worker, &jobs[t]);
    }
    for (int t = 0; t < threads; t++)
      pthread_join(th[t], NULL);
    free(jobs);
    free(th);
  }
  double t1 = now_s();

  /* octomap_mesh_masking.cpp:105-116: colorPassthrough, then the Octree constructor (updateNode per point,
   * updateInnerOccupancy, prune) on the concatenation of all frames. */
  orc_octree* tree = orc_octree_create(resolution);
  orc_point* filtered = (orc_point*)malloc(n * sizeof(orc_point));
  for (int f = 0; f < n_frames; f++)
  {
    const size_t m = orc_color_passthrough(annotated + (size_t)f * n, n, lower, upper, limit_negative, filtered);
    for (size_t i = 0; i < m; i++)
      res->inserted += (uint64_t)orc_octree_insert_point(tree, filtered[i].x, filtered[i].y, filtered[i].z);
  }
  orc_octree_update_inner_occupancy(tree);
  orc_octree_prune_tesseract(tree);
  double t2 = now_s();

  res->seconds_annotate = t1 - t0;
  res->seconds_integrate = t2 - t1;
  res->points = (uint64_t)n * (uint64_t)n_frames;
  res->leaves = orc_octree_num_leaves(tree);
  orc_octree_destroy(tree);
  free(filtered);
  free(annotated);
}
