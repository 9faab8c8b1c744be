/*
 * pcs_b200.h -- C ABI of libpcs_b200.so: the B200-native (sm_100a CUDA) annotate-and-integrate hot path of
 * swri-robotics/point_cloud_segmentation.
 *
 * This is the drop-in boundary.  Every entry point takes plain pointers, sizes and PODs; no PCL / OpenCV /
 * tesseract / torch type crosses it.  The C++ adapters under point_cloud_segmentation_b200/cpp/include keep the
 * reference's class and function signatures and call these symbols; INTEGRATION.md shows the binding.
 * Each declaration cites the reference interface it replaces (file:line relative to the reference root).
 *
 * Conventions
 *   - Every function returns a pcs_status (0 = ok).  Nothing throws across the ABI.  After a non-zero return
 *     pcs_last_error() gives a thread-local, human readable message.
 *   - The caller owns every buffer it passes.  The library owns device memory behind pcs_ctx / pcs_map handles.
 *   - A pcs_ctx is bound to one CUDA device and one stream and is not re-entrant; distinct contexts are
 *     independent (one per GPU).
 *   - Functions without a suffix take HOST pointers and are synchronous (results are valid on return), like
 *     the reference's functions.  `_dev` variants take DEVICE pointers, enqueue on the context's stream and
 *     return without synchronising.
 *   - There is no CPU fallback: every function needs a CUDA device and fails with PCS_ERR_CUDA without one.
 */
#ifndef PCS_B200_H
#define PCS_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define PCS_API __attribute__((visibility("default")))
#else
#define PCS_API
#endif

#define PCS_VERSION 100

typedef enum pcs_status
{
  PCS_OK = 0,
  PCS_ERR_BAD_ARG = 1,  /* null pointer, bad size, lower > upper, ... (the reference asserts, utils.h:103-105) */
  PCS_ERR_CUDA = 2,     /* a CUDA runtime call or kernel failed, or no device */
  PCS_ERR_CAPACITY = 3, /* voxel map or caller output buffer too small */
  PCS_ERR_COMM = 4      /* reserved for the multi-GPU exchange */
} pcs_status;

/* pcl::PointXYZRGB memory layout: 32 bytes, 16-byte aligned.  x,y,z,data[3] then b,g,r,a (aliasing `float rgb`)
 * then 12 bytes of padding.  Annotated clouds written by the library carry data[3] = 1.0f, a = 0 (the reference's
 * imageToCloud packs rgb into an int32, utils.h:161-162) and zero padding. */
typedef struct pcs_point_xyzrgb
{
  float x, y, z, data3;
  uint8_t b, g, r, a;
  uint32_t pad[3];
} pcs_point_xyzrgb;

/* Thresholds of hsvThresholdingDetector; defaults are the reference's compile-time constants
 * (pcs_detection/include/pcs_detection/hsv_thresholding.h:47-52).  Bounds are inclusive (cv::inRange).
 * morph_open = 1 adds a 5x5 MORPH_OPEN after thresholding (no counterpart in hsvThresholdingDetector; semantics of
 * pcs_detection/src_python/pcs_detection/models/histogram_backprojection.py:40-41).  Order: threshold, open, invert. */
typedef struct pcs_hsv_params
{
  int32_t h_lo, h_hi, s_lo, s_hi, v_lo, v_hi;
  int32_t inverted;
  int32_t morph_open;
} pcs_hsv_params;

/* Parameters of the fused annotate(+integrate) call.
 * lower/upper/limit_negative are colorPassthrough's (octomap_mesh_masking.h:50-53) applied to the annotation colour:
 * a point whose annotated r,g,b pass goes to count1 of its voxel (it is a point the reference would insert into the
 * octree), any other finite in-range point goes to count0.
 * world_frame_output = 1 writes pose-transformed coordinates into the annotated cloud (finite points only,
 * pcl::transformPointCloud semantics, octomap_mesh_mask_node.cpp:56-57); 0 keeps the sensor frame exactly as
 * imageToCloud does (utils.h:152-154). */
typedef struct pcs_annotate_params
{
  pcs_hsv_params hsv;
  int32_t lower, upper, limit_negative;
  int32_t world_frame_output;
} pcs_annotate_params;

/* MaskType values are the action's integers (pcs_msgs/action/ApplyOctomapMeshMask.action:19-21,
 * pcs_ros/src/octomap_mesh_mask_node.cpp:83-99) and OctomapMeshMask::MaskType's order (octomap_mesh_masking.h:65-70). */
typedef enum pcs_mask_type
{
  PCS_RETURN_INSIDE = 0,
  PCS_RETURN_OUTSIDE = 1,
  PCS_RETURN_COLORIZED = 2
} pcs_mask_type;

typedef struct pcs_stats
{
  uint64_t kernel_launches; /* kernels this context has launched so far */
  uint64_t points_in;       /* points seen by annotate / integrate calls */
  uint64_t bytes_h2d, bytes_d2h;
} pcs_stats;

/* Counters of a voxel map: what the scan has put into it so far. */
typedef struct pcs_map_stats
{
  uint64_t voxels;          /* distinct voxels touched */
  uint64_t occupied;        /* voxels with count1 >= 1: the leaves of the reference octree */
  uint64_t points_rejected; /* sum of count0: finite in-range points that failed colorPassthrough */
  uint64_t points_labelled; /* sum of count1: points the reference would have inserted into the octree */
  uint64_t table_slots;     /* size of the hash table */
  double load_factor;       /* voxels / table_slots */
} pcs_map_stats;

typedef struct pcs_ctx pcs_ctx;
typedef struct pcs_map pcs_map;

/* ---- context, errors, memory -------------------------------------------------------------------------------- */
PCS_API int pcs_version(void);
PCS_API const char* pcs_last_error(void);
/* `cuda_stream` is a cudaStream_t (may be NULL: the library creates its own non-blocking stream). */
PCS_API int pcs_ctx_create(int device, void* cuda_stream, pcs_ctx** ctx);
PCS_API int pcs_ctx_destroy(pcs_ctx* ctx);
PCS_API int pcs_ctx_sync(pcs_ctx* ctx);
PCS_API int pcs_get_stats(pcs_ctx* ctx, pcs_stats* out);
PCS_API int pcs_hsv_default_params(pcs_hsv_params* p);
PCS_API int pcs_annotate_default_params(pcs_annotate_params* p);
/* pinned host memory so that the host-pointer entry points can overlap copies with kernels */
PCS_API void* pcs_host_alloc(size_t bytes);
PCS_API void pcs_host_free(void* p);
/* device memory + copies + events for callers that keep data resident (benchmarks, pipelines) */
PCS_API int pcs_dev_alloc(pcs_ctx* ctx, size_t bytes, void** dptr);
PCS_API int pcs_dev_free(pcs_ctx* ctx, void* dptr);
PCS_API int pcs_memcpy_h2d(pcs_ctx* ctx, void* dst_dev, const void* src_host, size_t bytes);
PCS_API int pcs_memcpy_d2h(pcs_ctx* ctx, void* dst_host, const void* src_dev, size_t bytes);
PCS_API int pcs_memset_dev(pcs_ctx* ctx, void* dst_dev, int value, size_t bytes);
PCS_API int pcs_event_create(pcs_ctx* ctx, void** event);
PCS_API int pcs_event_record(pcs_ctx* ctx, void* event);
/* makes the context's stream wait for `event` (recorded by another context of the same process): lets a second context on
 * a side stream run the map exchange of one scan while the first integrates the next */
PCS_API int pcs_stream_wait_event(pcs_ctx* ctx, void* event);
PCS_API int pcs_event_elapsed_ms(void* start, void* stop, float* ms); /* synchronises on `stop` */
PCS_API int pcs_event_destroy(void* event);

/* ---- stage 1: 2D detector ------------------------------------------------------------------------------------- */
/* Replaces pcs_detection::hsvThresholdingDetector(const cv::Mat&, cv::Mat& mask, bool inverted)
 * (pcs_detection/include/pcs_detection/hsv_thresholding.h:45-70): 8UC3 BGR image -> 8UC1 mask in {0,255}.
 * `stride`/`mask_stride` are row pitches in bytes (cv::Mat::step). */
PCS_API int pcs_hsv_mask(pcs_ctx* ctx, const uint8_t* bgr, int width, int height, size_t stride,
                         const pcs_hsv_params* params, uint8_t* mask, size_t mask_stride);
PCS_API int pcs_hsv_mask_dev(pcs_ctx* ctx, const uint8_t* bgr_dev, int width, int height, size_t stride,
                             const pcs_hsv_params* params, uint8_t* mask_dev, size_t mask_stride);
/* Replaces the detector service body pcs_ros::process_image (pcs_ros/src/hsv_thresholding_node.cpp:8-28):
 * mask replicated to 3 channels (cv::merge, :20-23), 8UC3 in -> 8UC3 out, tightly packed rows. */
PCS_API int pcs_hsv_annotation_image(pcs_ctx* ctx, const uint8_t* bgr, int width, int height, size_t stride,
                                     const pcs_hsv_params* params, uint8_t* annotation_bgr);

/* Replaces HistogramBackprojection.annotate_image
 * (pcs_detection/src_python/pcs_detection/models/histogram_backprojection.py:22-52; service in
 * pcs_ros/src/histogram_backprojection_node:29-44): BGR2HSV, calcBackProject of the (hue, saturation) histogram
 * (h_bins x s_bins float32, row-major, ranges [0,180) x [0,255)), 5x5 ellipse filter2D, threshold (> threshold_min ->
 * threshold_value; the reference uses 30 and 150), 5x5 opening, min-max normalisation to {0, 255}, three equal channels.
 * 8UC3 in (row pitch `stride`) -> 8UC3 out, tightly packed. */
PCS_API int pcs_histogram_backprojection(pcs_ctx* ctx, const uint8_t* bgr, int width, int height, size_t stride,
                                         const float* histogram, int h_bins, int s_bins, int threshold_min,
                                         int threshold_value, uint8_t* annotation_bgr);

/* ---- stage 2: organized cloud <-> images, labelling, transform -------------------------------------------------- */
/* Replaces pcs_detection::cloudToImage (pcs_detection/include/pcs_detection/utils.h:98-129): organized cloud ->
 * CV_64FC3 position image (h*w*3 doubles) + CV_8UC3 BGR image (h*w*3 bytes).  1x1 clouds are rejected (:103). */
PCS_API int pcs_cloud_to_image(pcs_ctx* ctx, const pcs_point_xyzrgb* cloud, int width, int height,
                               double* position_image, uint8_t* bgr_image);
/* Replaces pcs_detection::imageToCloud (utils.h:138-171): 3-channel annotation + position image -> XYZRGB cloud whose
 * colour is the annotation. */
PCS_API int pcs_image_to_cloud(pcs_ctx* ctx, const uint8_t* annotation_bgr, const double* position_image, int width,
                               int height, pcs_point_xyzrgb* cloud_out);
/* Replaces pcs_scan_integration::colorPassthrough (octomap_mesh_masking.h:50-53, octomap_mesh_masking.cpp:52-96):
 * order-preserving compaction of the points whose r,g,b all pass.  `out` needs room for n points.
 * lower > upper -> PCS_ERR_BAD_ARG (the reference logs and asserts, .cpp:65-69). */
PCS_API int pcs_color_passthrough(pcs_ctx* ctx, const pcs_point_xyzrgb* in, uint64_t n, int lower, int upper,
                                  int limit_negative, pcs_point_xyzrgb* out, uint64_t* n_out);
/* No counterpart in the reference.  Same filter as pcs_color_passthrough, for hosts that must not let the library
 * allocate: a per-frame callback where cudaMalloc / cudaFree would stall every stream on the device, or a GPU shared
 * with a process that reserved nearly all device memory at start-up, where a late allocation can fail.  The call
 * itself allocates nothing (no cudaMalloc / cudaFree, no library scratch growth):
 *   - in_dev, out_dev (room for n points) and n_out_dev (one uint64_t) are DEVICE pointers owned by the caller; the
 *     work is enqueued on the context's stream and the count stays on the device;
 *   - the per-block counts and offsets of the compaction live in a fixed 48.25 KiB array inside the library image,
 *     which every GPU that loads the library holds until the process exits;
 *   - n may not exceed PCS_NOALLOC_MAX_POINTS (one 3840x2160 frame fits); more returns PCS_ERR_CAPACITY;
 *   - all calls on a device share that array and it keeps the previous call's contents (nothing in it needs to start
 *     at zero); calls whose kernels overlap on the GPU race on it, so keep them ordered;
 *   - in_dev and out_dev must be distinct, non-overlapping ranges (the filter cannot run in place; an overlap is
 *     refused with PCS_ERR_BAD_ARG); n_out_dev must be 8-byte aligned;
 *   - the first call on a device looks up the array's device address once, and with CUDA's lazy module loading it is
 *     also the call at which the driver loads the library's device code (the only point where memory may be reserved on
 *     the caller's behalf): make that first call early, before the time-critical phase and before any stream capture. */
#define PCS_NOALLOC_MAX_POINTS (1ull << 23)
PCS_API int pcs_color_passthrough_noalloc_dev(pcs_ctx* ctx, const pcs_point_xyzrgb* in_dev, uint64_t n, int lower,
                                              int upper, int limit_negative, pcs_point_xyzrgb* out_dev,
                                              uint64_t* n_out_dev);
/* Replaces pcl_ros::transformPointCloud as called at pcs_ros/src/octomap_mesh_mask_node.cpp:56-57: `pose` is a
 * row-major float 4x4; x' = ((m00 x + m01 y) + m02 z) + m03 in float without contraction; non-finite points are
 * copied unchanged. */
PCS_API int pcs_transform_cloud(pcs_ctx* ctx, const float* pose, const pcs_point_xyzrgb* in, uint64_t n,
                                pcs_point_xyzrgb* out);

/* Fused replacement of one PointCloudAnnotator batch with the HSV detector plugged in:
 *   addPointCloud -> cloudToImage   (pcs_detection/src/pcs_detection/point_cloud_annotator.cpp:7-30)
 *   annotateImages -> detector -> imageToCloud per frame -> results (point_cloud_annotator.cpp:32-82,
 *   pcs_ros/src/point_cloud_annotator_node.cpp:68-99, pcs_ros/src/hsv_thresholding_node.cpp:8-28)
 * and, when `map` is not NULL, of the integration of every annotated frame into the voxel map
 * (OctomapMeshMask::setOctree, octomap_mesh_masking.cpp:105-116, on the concatenation of the frames).
 *   clouds        n_frames * width * height points, frame-major, organized row-major
 *   poses         n_frames row-major float 4x4 camera->world, or NULL for identity
 *   annotated_out n_frames * width * height points (may be NULL: integrate-only)
 *   mask_out      n_frames * width * height bytes in {0,255} (may be NULL)
 * hsv.morph_open = 1 runs the detector as a separate image kernel first. */
PCS_API int pcs_annotate_frames(pcs_ctx* ctx, pcs_map* map, const pcs_point_xyzrgb* clouds, int n_frames, int width,
                                int height, const float* poses, const pcs_annotate_params* params,
                                pcs_point_xyzrgb* annotated_out, uint8_t* mask_out);
PCS_API int pcs_annotate_frames_dev(pcs_ctx* ctx, pcs_map* map, const pcs_point_xyzrgb* clouds_dev, int n_frames,
                                    int width, int height, const float* poses_dev,
                                    const pcs_annotate_params* params, pcs_point_xyzrgb* annotated_out_dev,
                                    uint8_t* mask_out_dev);

/* The compacted product of stage 2: the labelled points the reference hands to the octree, i.e. what
 * colorPassthrough(transformPointCloud(annotated frame)) keeps (octomap_mesh_mask_node.cpp:56-68,
 * octomap_mesh_masking.cpp:110-115) minus the points octomap would silently skip (non-finite, or outside the key range
 * of `resolution`; pass resolution = 0 to skip the range test).  World-frame coordinates, colour = annotation, input
 * order preserved (frame-major, row-major).  out == NULL queries the count. */
PCS_API int pcs_labelled_points(pcs_ctx* ctx, const pcs_point_xyzrgb* clouds, int n_frames, int width, int height,
                                const float* poses, const pcs_annotate_params* params, double resolution, uint64_t cap,
                                pcs_point_xyzrgb* out, uint64_t* n_out);

/* ---- stage 3: voxel map (stands in for octomap::OcTree behind tesseract_geometry::Octree) ------------------------- */
/* A map is a device hash of 48-bit voxel keys (key48 = kz<<32 | ky<<16 | kx, each k = floor(coord / resolution) +
 * 32768 as octomap's coordToKeyChecked) with two hit counters per voxel.  The reference octree for the filter used
 * at integration time is exactly { key : count1 >= 1 } with log-odds LUT[min(count1, 5)]
 * (octomap hit = logodds(0.7), clamp logodds(0.971)).  `capacity_voxels` bounds the number of distinct voxels: the table
 * has at least 2 x capacity_voxels slots, and pcs_map_size / pcs_map_export report PCS_ERR_CAPACITY as soon as more than
 * capacity_voxels distinct voxels have been inserted (results stay exact up to the slot count, but buffers sized from
 * capacity_voxels would no longer fit).
 * Limit: the two counters of a voxel are the halves of ONE 64-bit word updated by a single atomic add, so count0 is exact
 * only below 2^32 rejected points per voxel -- at 2^32 it would carry into count1.  A voxel of the BASELINE scans sees
 * ~10^4 points; a long-lived accumulating map must be exported and cleared before any voxel can reach 4.29e9 points. */
PCS_API int pcs_map_create(pcs_ctx* ctx, double resolution, uint64_t capacity_voxels, pcs_map** map);
PCS_API int pcs_map_destroy(pcs_ctx* ctx, pcs_map* map);
PCS_API int pcs_map_clear(pcs_ctx* ctx, pcs_map* map);
/* Replaces OctomapMeshMask::setOctree(cloud, resolution, lower, upper, limit_negative)
 * (octomap_mesh_masking.h:105-109, octomap_mesh_masking.cpp:105-116) in accumulate form: colorPassthrough on the
 * points' own r,g,b, then one hit per finite in-range point.  `pose` may be NULL. */
PCS_API int pcs_map_integrate(pcs_ctx* ctx, pcs_map* map, const pcs_point_xyzrgb* points, uint64_t n,
                              const float* pose, int lower, int upper, int limit_negative);
PCS_API int pcs_map_integrate_dev(pcs_ctx* ctx, pcs_map* map, const pcs_point_xyzrgb* points_dev, uint64_t n,
                                  const float* pose_dev, int lower, int upper, int limit_negative);
/* number of voxels touched / occupied (count1 >= 1); PCS_ERR_CAPACITY if the table overflowed at any point */
PCS_API int pcs_map_size(pcs_ctx* ctx, pcs_map* map, uint64_t* n_voxels, uint64_t* n_occupied);
/* finite / labelled point counts, voxels touched and occupied, hash load factor (one pass over the live voxels) */
PCS_API int pcs_map_get_stats(pcs_ctx* ctx, pcs_map* map, pcs_map_stats* out);
/* Export in octree (Morton / child-index DFS) order: key48, the two counters and the float log-odds the reference
 * tree would hold (0 for voxels with count1 == 0).  Any output pointer may be NULL.  If cap < n only `n` is
 * reported and PCS_ERR_CAPACITY returned (two-call size query). */
PCS_API int pcs_map_export(pcs_ctx* ctx, pcs_map* map, uint64_t cap, uint64_t* key48, uint32_t* count0,
                           uint32_t* count1, float* logodds, uint64_t* n);
/* OcTree-compatible views of the map.
 * pcs_map_write_bt: the occupied voxels (count1 >= 1) as an octomap binary tree, byte-compatible with
 * octomap::OcTree::writeBinary for the same occupied set -- the format the reference's test reads with
 * octomap::OcTree(path) (pcs_scan_integration/test/octomap_mesh_masking_unit.cpp:140-141).  out == NULL queries the size.
 * pcs_map_leaf_centers: one point per voxel at the voxel centre, in octree order, r = g = b = 255 where count1 >= 1 else
 * 0 (alpha 255): the colourised centre cloud the mesh-mask action consumes
 * (pcs_ros/src/octomap_mesh_mask_node.cpp:47-68).  points == NULL queries the count. */
PCS_API int pcs_map_write_bt(pcs_ctx* ctx, pcs_map* map, uint8_t* out, uint64_t cap, uint64_t* n_bytes);
PCS_API int pcs_map_leaf_centers(pcs_ctx* ctx, pcs_map* map, uint64_t cap, pcs_point_xyzrgb* points, uint64_t* n_out);
/* Import / merge records (checkpoint restore, multi-GPU merge): counters are added to existing voxels. */
PCS_API int pcs_map_merge(pcs_ctx* ctx, pcs_map* map, const uint64_t* key48, const uint32_t* count0,
                          const uint32_t* count1, uint64_t n);
/* Multi-GPU exchange support: pack every voxel as a 16-byte record {key48, count0 | count1 << 32} grouped by owner
 * = mix64(key48) % n_owners into `records_dev` (room for pcs_map_size voxels); owner_counts (host, n_owners) receives
 * the group sizes.  pcs_map_merge_records_dev adds such records (e.g. received from peers) into a map. */
PCS_API int pcs_map_pack_by_owner_dev(pcs_ctx* ctx, pcs_map* map, int n_owners, void* records_dev,
                                      uint64_t cap_records, uint64_t* owner_counts);
PCS_API int pcs_map_merge_records_dev(pcs_ctx* ctx, pcs_map* map, const void* records_dev, uint64_t n);
PCS_API uint64_t pcs_mix64(uint64_t key48);
/* Peer-memory form of the exchange (one process per GPU, NVLink / NVSwitch): each rank exports its OWNED map with
 * pcs_map_ipc_export (128 bytes holding the CUDA IPC handle, which the host exchanges by any means), opens every peer's
 * handle with pcs_map_ipc_open (same resolution / capacity as at creation; destroy with pcs_map_destroy), and then one
 * kernel, pcs_map_push_to_owners_dev, adds every voxel of the LOCAL map directly into owners[mix64(key48) % n_owners]
 * with system-scope atomics -- no packing, no staging buffers, no collective for the data.  clear_local = 1 leaves the
 * local map empty (as pcs_map_clear would), saving that launch when the next scan follows.  The host must put a
 * barrier across ranks before the push (owned maps cleared everywhere) and after it (all pushes landed).
 * Owner maps may equally be ordinary maps of this process (single-GPU use and tests). */
PCS_API int pcs_map_ipc_export(pcs_ctx* ctx, pcs_map* map, void* handle128);
PCS_API int pcs_map_ipc_open(pcs_ctx* ctx, const void* handle128, double resolution, uint64_t capacity_voxels,
                             pcs_map** peer_map);
PCS_API int pcs_map_push_to_owners_dev(pcs_ctx* ctx, pcs_map* local, pcs_map* const* owners, int n_owners,
                                       int clear_local);
/* Mailbox form of the same exchange (default of distributed.PeerMapExchange): nothing atomic crosses NVLink.  Every rank
 * owns, per owned map, an INBOX of pcs_map_inbox_bytes(capacity, n) bytes (a 256-byte header of per-source record counts,
 * then one segment of `seg_capacity` 16-byte records per source rank) that its peers map through pcs_ipc_export_dev /
 * pcs_ipc_open_dev.  pcs_map_push_mailbox_dev adds the voxels this rank owns to `own` with local atomics and stores every
 * other voxel as a record into this rank's segment of its owner's inbox (peer_inboxes[o], o != rank; posted stores, slots
 * from a local cursor), then publishes the counts.  After the barrier pcs_map_merge_inbox_dev adds the records of the own
 * inbox to `own`.  An inbox may be written again only after its owner has merged it: alternate two, like the owned maps. */
PCS_API uint64_t pcs_map_inbox_bytes(uint64_t capacity_voxels, int n_owners);
PCS_API int pcs_ipc_export_dev(pcs_ctx* ctx, void* dptr, void* handle64);
PCS_API int pcs_ipc_open_dev(pcs_ctx* ctx, const void* handle64, void** dptr);
PCS_API int pcs_ipc_close_dev(pcs_ctx* ctx, void* dptr);
PCS_API int pcs_map_push_mailbox_dev(pcs_ctx* ctx, pcs_map* local, pcs_map* own, int rank, int n_owners,
                                     void* const* peer_inboxes, uint64_t seg_capacity, int clear_local);
PCS_API int pcs_map_merge_inbox_dev(pcs_ctx* ctx, pcs_map* own, const void* inbox, int rank, int n_owners,
                                    uint64_t seg_capacity);
/* The barrier of that exchange without a collective (no counterpart in the reference, which is single-process): every map
 * carries 16 epoch words that clears do not touch.  After its push, rank r writes the epoch number into word r of every
 * peer's owned map with pcs_stream_write_u32_dev (cuStreamWriteValue32: ordered after the push in the context's stream)
 * and waits with pcs_stream_wait_geq_u32_dev (cuStreamWaitValue32, >=) until the words of all peers in its own map have
 * reached that epoch.  Nothing spins on an SM and the host does not block.  pcs_map_epoch_flag_dev returns the device
 * address of word `index` of a local or IPC-opened map. */
PCS_API int pcs_map_epoch_flag_dev(pcs_map* map, int index, void** flag_dev);
PCS_API int pcs_stream_write_u32_dev(pcs_ctx* ctx, void* flag_dev, uint32_t value);
PCS_API int pcs_stream_wait_geq_u32_dev(pcs_ctx* ctx, void* flag_dev, uint32_t value);
/* The same barrier as ONE 32-thread kernel: lane i stores `epoch` into their_flags[i] (this rank's word in peer i's map,
 * release at system scope) and then waits until my_flags[i] (peer i's word in this rank's map) has reached it.  One launch
 * instead of 2 n stream operations.  One process per GPU is assumed (the waiting kernel and the kernels it waits for never
 * share a device); a peer that does not arrive within ~10 s makes the kernel trap rather than hang. */
PCS_API int pcs_exchange_barrier_dev(pcs_ctx* ctx, void* const* their_flags, void* const* my_flags, int n, uint32_t epoch);

/* ---- stage 4: mesh mask -------------------------------------------------------------------------------------------- */
/* Replaces OctomapMeshMask::maskMesh(mask_type) (octomap_mesh_masking.h:131, octomap_mesh_masking.cpp:118-283) against
 * the occupied voxels (count1 >= 1) of `map`.
 *   verts   nv * 3 doubles (tesseract_common::VectorVector3d);  faces4: nt * [3, i0, i1, i2] (Eigen::VectorXi layout)
 *   a triangle is "inside" iff it touches or penetrates an occupied voxel box (closed test, contact distance <= 0)
 *   INSIDE / OUTSIDE: un-welded output in input order, 3 vertices + face [3, n-3, n-2, n-1] per kept triangle;
 *     INSIDE lists each contacted triangle once (the reference lists it once per Bullet contact, .cpp:196-216)
 *   COLORIZED: input mesh copied, out_rgb (255,0,0) on vertices of contacted triangles else (0,128,0) (.cpp:166-181)
 *   out_verts / out_faces4 / out_rgb may be NULL to query sizes; out_cap_tris = room in triangles
 *     (out_verts needs 3 * out_cap_tris vertices for INSIDE/OUTSIDE, nv for COLORIZED)
 *   tri_flags (nt bytes, may be NULL) receives 1 for contacted triangles */
PCS_API int pcs_mesh_mask(pcs_ctx* ctx, pcs_map* map, const double* verts, uint64_t nv, const int32_t* faces4,
                          uint64_t nt, int mask_type, uint64_t out_cap_tris, double* out_verts, int32_t* out_faces4,
                          uint8_t* out_rgb, uint8_t* tri_flags, uint64_t* n_out_verts, uint64_t* n_out_tris);
/* Device-resident forms (mesh and outputs stay in HBM, work is enqueued on the context's stream).
 * pcs_mesh_mask_dev: the whole of maskMesh -- contact flags, then the order-preserving compaction that emits the kept
 *   triangles (INSIDE / OUTSIDE: 3 un-welded vertices, one face [3, 3i, 3i+1, 3i+2] and 9 colour bytes per kept triangle,
 *   up to out_cap_tris of them; the kept count, which may exceed out_cap_tris, lands in *n_out_tris_dev) or the per-vertex
 *   colours (COLORIZED: out_rgb_dev holds nv * 3 bytes, vertices / faces are the input's and are not copied).  The output
 *   pointers may be NULL to count only.  tri_flags_dev (nt bytes) is required.
 * pcs_mesh_contacts_dev: the contact flags only (1 byte per triangle). */
PCS_API int pcs_mesh_mask_dev(pcs_ctx* ctx, pcs_map* map, const double* verts_dev, uint64_t nv, const int32_t* faces4_dev,
                              uint64_t nt, int mask_type, uint64_t out_cap_tris, double* out_verts_dev,
                              int32_t* out_faces4_dev, uint8_t* out_rgb_dev, uint8_t* tri_flags_dev,
                              uint64_t* n_out_tris_dev);
/* The output assembly alone, from contact flags that already exist (the multi-GPU mask: every rank decides a range of
 * triangles, the flags are gathered, one rank emits): same outputs as pcs_mesh_mask_dev.  *n_out_tris_dev receives the
 * kept count of the compaction (INSIDE / OUTSIDE); it is set to 0 for COLORIZED, which keeps all nt triangles. */
PCS_API int pcs_mesh_emit_dev(pcs_ctx* ctx, const double* verts_dev, uint64_t nv, const int32_t* faces4_dev, uint64_t nt,
                              const uint8_t* tri_flags_dev, int mask_type, uint64_t out_cap_tris, double* out_verts_dev,
                              int32_t* out_faces4_dev, uint8_t* out_rgb_dev, uint64_t* n_out_tris_dev);
PCS_API int pcs_mesh_contacts_dev(pcs_ctx* ctx, pcs_map* map, const double* verts_dev, uint64_t nv,
                                  const int32_t* faces4_dev, uint64_t nt, uint8_t* tri_flags_dev);

/* ---- synthetic scan generator (test / benchmark input) ---------------------------------------------------- */
/* Fills n_frames organized frames starting at `first_frame`; poses_dev holds n_frames row-major 4x4 floats.
 * Bit-identical to oracle/synth.c. */
PCS_API int pcs_synth_frames_dev(pcs_ctx* ctx, uint64_t seed, int first_frame, int n_frames, int width, int height,
                                 const float* poses_dev, pcs_point_xyzrgb* out_dev);

#ifdef __cplusplus
}
#endif
#endif /* PCS_B200_H */
