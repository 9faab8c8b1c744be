// annotate.cu -- stage 2 (+3): organized-cloud labelling fused with the detector and the voxel-map update.
//
// One streaming pass per point replaces, for a batch of frames,
//   cloudToImage            pcs_detection/include/pcs_detection/utils.h:98-129       (no position / BGR image is built)
//   hsvThresholdingDetector pcs_detection/include/pcs_detection/hsv_thresholding.h:45-70
//   cv::merge(mask x3)      pcs_ros/src/hsv_thresholding_node.cpp:20-23
//   imageToCloud            pcs_detection/include/pcs_detection/utils.h:138-171
//   transformPointCloud     pcs_ros/src/octomap_mesh_mask_node.cpp:56-57
//   colorPassthrough        pcs_scan_integration/src/pcs_scan_integration/octomap_mesh_masking.cpp:52-96
//   Octree(cloud, res)      octomap_mesh_masking.cpp:114-115 (octomap updateNode per point -> hash of hit counts)
//
// Data movement per point (HBM-bound): 32 B PointXYZRGB read with one 256-bit load, 32 B annotated
// PointXYZRGB written with one 256-bit store, 1 B mask written -> 65 B; the voxel table is small and L2-resident and
// is touched through warp-aggregated 64-bit atomics (one per distinct voxel per warp).
#include "common.cuh"
#include "compact.cuh"
#include "annotate.cuh"
#include "hsv.cuh"

namespace pcs
{
#ifndef PCS_ANN_UNROLL
#define PCS_ANN_UNROLL 1
#endif
#ifndef PCS_ANN_ROUNDS
#define PCS_ANN_ROUNDS 4
#endif
#ifndef PCS_ANN_MIN_BLOCKS
#define PCS_ANN_MIN_BLOCKS 6
#endif
#ifndef PCS_ANN_ROUND_UNROLL
#define PCS_ANN_ROUND_UNROLL 1
#endif
#ifndef PCS_ANN_PERSISTENT
#define PCS_ANN_PERSISTENT 0
#endif
#ifndef PCS_ANN_THREADS
#define PCS_ANN_THREADS 256
#endif
constexpr int kAnnThreads = PCS_ANN_THREADS;
constexpr int kAnnRoundUnroll = PCS_ANN_ROUND_UNROLL;
constexpr int kAnnUnroll = PCS_ANN_UNROLL;            // points in flight per thread
constexpr int kAnnRounds = PCS_ANN_ROUNDS;            // rounds per tile
constexpr int kAnnTile = kAnnThreads * kAnnUnroll * kAnnRounds;  // 1024 points: one pose / frame lookup per tile

template <bool HAS_MAP, bool WRITE_ANN, bool WRITE_MASK>
__global__ void __launch_bounds__(kAnnThreads, PCS_ANN_MIN_BLOCKS) annotate_kernel(const AnnotateArgs a)
{
  __shared__ int s_lut[512];
  bool lut_ready = false;
  const bool has_pose = (a.poses != nullptr);
  // One tile per CTA by default (grid = n_tiles): measured on B200, a plain 32 B/point copy with this tile shape moves
  // 6.8 TB/s when the hardware block scheduler hands out the tiles and only 6.1 TB/s from 592 persistent CTAs that
  // stride over them (tools/copy_probe.cu), so the loop below normally runs once.
  for (unsigned int tile = blockIdx.x; tile < a.n_tiles; tile += gridDim.x)
  {
    const unsigned int frame = tile / a.tiles_per_frame;
    const unsigned int t_in_frame = tile - frame * a.tiles_per_frame;
    const unsigned long long frame_base = (unsigned long long)frame * a.pts_per_frame;
    Pose P;
    if (has_pose && (HAS_MAP || WRITE_ANN))
    {
      const float4* src = reinterpret_cast<const float4*>(a.poses + (size_t)frame * 16);
#pragma unroll
      for (int i = 0; i < 3; i++)
      {
        const float4 v = __ldg(src + i);
        P.m[4 * i + 0] = v.x;
        P.m[4 * i + 1] = v.y;
        P.m[4 * i + 2] = v.z;
        P.m[4 * i + 3] = v.w;
      }
    }
    // Software pipeline over the rounds of a tile:
    //   issue the 256-bit loads of round r + 1  ->  process round r (detector, transform, key, warp aggregation,
    //   voxel update).  Measured on B200: splitting the voxel update into "issue the home-slot load now, add one
    //   round later" was slower (80 vs 89 G points/s), as was more occupancy at fewer registers; unrolling the round
    //   loop changes nothing, so it stays rolled and the body stays small.
    unsigned int local = t_in_frame * kAnnTile + threadIdx.x;
    const Point8* src_pts = a.cloud + frame_base;
    Point8 cur[kAnnUnroll], nxt[kAnnUnroll];
#pragma unroll
    for (int u = 0; u < kAnnUnroll; u++)
      if (local + u * kAnnThreads < a.pts_per_frame)
        cur[u] = ld_point_stream(src_pts + local + u * kAnnThreads);
    if (!lut_ready)
    {
      // the detector's tables (2 KB, L2-resident) are fetched while the first points are already in flight
      load_hsv_lut(s_lut);
      __syncthreads();
      lut_ready = true;
    }
#pragma unroll kAnnRoundUnroll
    for (int r = 0; r < kAnnRounds; r++)
    {
      const unsigned int next_local = local + kAnnUnroll * kAnnThreads;
      if (r + 1 < kAnnRounds)
      {
#pragma unroll
        for (int u = 0; u < kAnnUnroll; u++)
          if (next_local + u * kAnnThreads < a.pts_per_frame)
            nxt[u] = ld_point_stream(src_pts + next_local + u * kAnnThreads);
      }
#pragma unroll
      for (int u = 0; u < kAnnUnroll; u++)
      {
        const unsigned int l = local + u * kAnnThreads;
        annotate_point<HAS_MAP, WRITE_ANN, WRITE_MASK>(a, s_lut, P, has_pose, cur[u], frame_base + l, l < a.pts_per_frame);
      }
#pragma unroll
      for (int u = 0; u < kAnnUnroll; u++)
        cur[u] = nxt[u];
      local = next_local;
    }
  }
}

// ---- plain point list into the map (setOctree on an arbitrary coloured cloud) ------------------------------------------
__global__ void __launch_bounds__(256, 4) integrate_kernel(const Point8* __restrict__ pts, unsigned long long n,
                                                          const float* __restrict__ pose, int lower, int upper, int neg,
                                                          MapView map)
{
  float m[12];
  const bool has_pose = pose != nullptr;
  if (has_pose)
  {
#pragma unroll
    for (int i = 0; i < 12; i++)
      m[i] = __ldg(pose + i);
  }
  const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
  const unsigned long long n_round = (n + 31ull) & ~31ull;  // whole warps stay converged for the ballots
  for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_round; i += stride)
  {
    const bool in = i < n;
    Point8 p;
    p.x = p.y = p.z = 0.f;
    p.rgba = 0;
    if (in)
      p = ld_point_stream(pts + i);
    float wx = p.x, wy = p.y, wz = p.z;
    if (has_pose)
    {
      wx = xform_row(m[0], m[1], m[2], m[3], p.x, p.y, p.z);
      wy = xform_row(m[4], m[5], m[6], m[7], p.x, p.y, p.z);
      wz = xform_row(m[8], m[9], m[10], m[11], p.x, p.y, p.z);
    }
    unsigned int kx, ky, kz;
    bool valid = in && finite3(wx, wy, wz);  // a non-finite input stays non-finite through the transform
    valid = coord_to_key(map.inv_res, wx, kx) && valid;
    valid = coord_to_key(map.inv_res, wy, ky) && valid;
    valid = coord_to_key(map.inv_res, wz, kz) && valid;
    const unsigned long long key = ((unsigned long long)kz << 32) | ((unsigned long long)ky << 16) | kx;
    const int b = p.rgba & 0xff, g = (p.rgba >> 8) & 0xff, r = (p.rgba >> 16) & 0xff;
    map_add_warp(map, valid, key, color_pass(r, g, b, lower, upper, neg));
  }
}

// ---- cloudToImage / imageToCloud / transform as stand-alone kernels (drop-in API, not on the fused path) ---------------
__global__ void cloud_to_image_kernel(const Point8* __restrict__ cloud, unsigned long long n, double* __restrict__ pos,
                                      uint8_t* __restrict__ bgr)
{
  const unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n)
    return;
  const Point8 p = ld_point_stream(cloud + i);
  if (pos)
  {
    pos[3 * i + 0] = (double)p.x;  // utils.h:117-119
    pos[3 * i + 1] = (double)p.y;
    pos[3 * i + 2] = (double)p.z;
  }
  if (bgr)
  {
    bgr[3 * i + 0] = (uint8_t)(p.rgba & 0xff);  // utils.h:122-124: B, G, R
    bgr[3 * i + 1] = (uint8_t)((p.rgba >> 8) & 0xff);
    bgr[3 * i + 2] = (uint8_t)((p.rgba >> 16) & 0xff);
  }
}

__global__ void image_to_cloud_kernel(const uint8_t* __restrict__ ann, const double* __restrict__ pos,
                                      unsigned long long n, Point8* __restrict__ out)
{
  const unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n)
    return;
  Point8 o;
  o.x = (float)pos[3 * i + 0];  // utils.h:152-154
  o.y = (float)pos[3 * i + 1];
  o.z = (float)pos[3 * i + 2];
  o.w = 1.0f;
  const unsigned int b = ann[3 * i + 0], g = ann[3 * i + 1], r = ann[3 * i + 2];
  o.rgba = (r << 16) | (g << 8) | b;  // utils.h:161
  o.p0 = o.p1 = o.p2 = 0u;
  st_point_stream(out + i, o);
}

__global__ void transform_kernel(const Point8* __restrict__ in, unsigned long long n, const float* __restrict__ pose,
                                 Point8* __restrict__ out)
{
  const unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n)
    return;
  Point8 p = ld_point_stream(in + i);
  if (finite3(p.x, p.y, p.z))
  {
    const float x = p.x, y = p.y, z = p.z;
    p.x = xform_row(__ldg(pose + 0), __ldg(pose + 1), __ldg(pose + 2), __ldg(pose + 3), x, y, z);
    p.y = xform_row(__ldg(pose + 4), __ldg(pose + 5), __ldg(pose + 6), __ldg(pose + 7), x, y, z);
    p.z = xform_row(__ldg(pose + 8), __ldg(pose + 9), __ldg(pose + 10), __ldg(pose + 11), x, y, z);
  }
  st_point_stream(out + i, p);
}

struct PassPred
{
  const Point8* pts;
  int lower, upper, neg;
  __device__ bool operator()(unsigned long long i) const
  {
    const unsigned int c = reinterpret_cast<const unsigned int*>(pts + i)[4];
    return color_pass((c >> 16) & 0xff, (c >> 8) & 0xff, c & 0xff, lower, upper, neg);
  }
};

struct PassEmit
{
  const Point8* in;
  Point8* out;
  __device__ void operator()(unsigned long long i, unsigned long long dst) const
  {
    const uint4* s = reinterpret_cast<const uint4*>(in + i);
    uint4* d = reinterpret_cast<uint4*>(out + dst);
    d[0] = s[0];
    d[1] = s[1];
  }
};

// labelled-point list: finite after the transform, inside the voxel key range, annotation colour passes the filter
struct LabelledPred
{
  const Point8* pts;  // annotated, world frame
  double inv_res;     // 0: skip the key-range test
  int lower, upper, neg;
  __device__ bool operator()(unsigned long long i) const
  {
    const uint4 a = reinterpret_cast<const uint4*>(pts + i)[0];
    const unsigned int c = reinterpret_cast<const unsigned int*>(pts + i)[4];
    const float x = __uint_as_float(a.x), y = __uint_as_float(a.y), z = __uint_as_float(a.z);
    bool ok = finite3(x, y, z) && color_pass((c >> 16) & 0xff, (c >> 8) & 0xff, c & 0xff, lower, upper, neg);
    if (inv_res > 0.0)
    {
      unsigned int k;
      ok = coord_to_key(inv_res, x, k) && ok;
      ok = coord_to_key(inv_res, y, k) && ok;
      ok = coord_to_key(inv_res, z, k) && ok;
    }
    return ok;
  }
};

int fill_annotate_args(pcs_map* map, const Point8* clouds, unsigned int n_frames, unsigned int width, unsigned int height,
                       const float* poses, const pcs_annotate_params* prm, Point8* ann, uint8_t* mask, AnnotateArgs& a)
{
  const unsigned int pts_per_frame = width * height;
  a.cloud = clouds;
  a.ann = ann;
  a.mask = mask;
  a.poses = poses;
  a.pts_per_frame = pts_per_frame;
  a.width = width;
  a.height = height;
  a.tiles_per_frame = (pts_per_frame + kAnnTile - 1) / kAnnTile;
  if ((unsigned long long)a.tiles_per_frame * n_frames > 0x7fffffffull)
  {
    set_error("batch too large for one launch: %llu tiles (limit 2^31 - 1); split the batch",
              (unsigned long long)a.tiles_per_frame * n_frames);
    return PCS_ERR_BAD_ARG;
  }
  a.n_tiles = a.tiles_per_frame * n_frames;
  a.hsv = HsvBounds{ prm->hsv.h_lo, prm->hsv.h_hi, prm->hsv.s_lo, prm->hsv.s_hi, prm->hsv.v_lo, prm->hsv.v_hi };
  a.hsv_tables = nullptr;
  a.inverted = prm->hsv.inverted ? 1 : 0;
  // colorPassthrough on the annotation colours (0,0,0) and (255,255,255), octomap_mesh_masking.cpp:75-87
  const int lo = prm->lower, up = prm->upper;
  a.pass_black = prm->limit_negative ? (0 <= lo || 0 >= up) : (0 >= lo && 0 <= up);
  a.pass_white = prm->limit_negative ? (255 <= lo || 255 >= up) : (255 >= lo && 255 <= up);
  a.world_out = prm->world_frame_output ? 1 : 0;
  if (map)
    a.map = map->view;
  else
    a.map = MapView{ nullptr, 0, nullptr, nullptr, nullptr, 0.0 };
  return PCS_OK;
}

static int launch_annotate(pcs_ctx* ctx, const AnnotateArgs& a, bool hm, bool wa, bool wm)
{
#if PCS_ANN_PERSISTENT
  const unsigned int max_blocks = (unsigned int)ctx->sm_count * (unsigned int)PCS_ANN_MIN_BLOCKS;  // one resident wave
  const unsigned int grid = a.n_tiles < max_blocks ? a.n_tiles : max_blocks;
#else
  const unsigned int grid = a.n_tiles;  // one 1024-point tile per CTA
#endif
#define PCS_ANN_CASE(HM, WA, WM)                                                                                       \
  if (hm == HM && wa == WA && wm == WM)                                                                                \
    annotate_kernel<HM, WA, WM><<<grid, kAnnThreads, 0, ctx->stream>>>(a);
  PCS_ANN_CASE(true, true, true)
  PCS_ANN_CASE(true, true, false)
  PCS_ANN_CASE(true, false, true)
  PCS_ANN_CASE(true, false, false)
  PCS_ANN_CASE(false, true, true)
  PCS_ANN_CASE(false, true, false)
  PCS_ANN_CASE(false, false, true)
#undef PCS_ANN_CASE
  count_launch(ctx);
  return check_launch("annotate_kernel");
}

// One launch for the whole batch, with or without the opening: the detector runs inside the kernel either way (no BGR
// plane, no mask image and no per-frame launch).
int annotate_dev(pcs_ctx* ctx, pcs_map* map, const pcs_point_xyzrgb* clouds, int n_frames, int width, int height,
                 const float* poses, const pcs_annotate_params* prm, pcs_point_xyzrgb* ann, uint8_t* mask)
{
  const unsigned long long ppf = (unsigned long long)width * height;
  const bool hm = map != nullptr, wa = ann != nullptr, wm = mask != nullptr;
  if (!hm && !wa && !wm)
    return PCS_OK;  // nothing to produce
  AnnotateArgs a;
  PCS_TRY(fill_annotate_args(map, (const Point8*)clouds, (unsigned int)n_frames, (unsigned int)width, (unsigned int)height,
                             poses, prm, (Point8*)ann, mask, a));
  ctx->stats.points_in += ppf * n_frames;
  if (a.n_tiles == 0)
    return PCS_OK;
  if (prm->hsv.morph_open)
  {
    PCS_TRY(hsv_decision_tables(ctx, prm->hsv, &a.hsv_tables));
    if (!a.hsv_tables)
    {
      set_error("HSV thresholds cannot be tabulated");  // cannot happen for interval bounds; kept as a guard
      return PCS_ERR_BAD_ARG;
    }
    return launch_annotate_open(ctx, a, (unsigned int)n_frames, hm, wa, wm);
  }
  return launch_annotate(ctx, a, hm, wa, wm);
}
}  // namespace pcs

using namespace pcs;

extern "C" {

int pcs_annotate_frames_dev(pcs_ctx* ctx, pcs_map* map, const pcs_point_xyzrgb* clouds_dev, int n_frames, int width,
                            int height, const float* poses_dev, const pcs_annotate_params* params,
                            pcs_point_xyzrgb* annotated_out_dev, uint8_t* mask_out_dev)
{
  PCS_TRY(ctx_bind(ctx));
  PCS_CHECK_ARG(clouds_dev && params, "NULL argument");
  PCS_CHECK_ARG(n_frames >= 0 && width > 0 && height > 0, "bad frame shape");
  PCS_CHECK_ARG((unsigned long long)width * height < (1ull << 31), "frame too large");
  PCS_CHECK_ARG(params->lower <= params->upper, "colorPassthrough: lower limit greater than upper limit");
  PCS_CHECK_ARG(((uintptr_t)clouds_dev & 31u) == 0 && ((uintptr_t)annotated_out_dev & 31u) == 0,
                "point buffers must be 32-byte aligned");
  PCS_CHECK_ARG(((uintptr_t)poses_dev & 15u) == 0, "poses must be 16-byte aligned");
  if (n_frames == 0)
    return PCS_OK;
  return annotate_dev(ctx, map, clouds_dev, n_frames, width, height, poses_dev, params, annotated_out_dev, mask_out_dev);
}

int pcs_annotate_frames(pcs_ctx* ctx, pcs_map* map, const pcs_point_xyzrgb* clouds, int n_frames, int width, int height,
                        const float* poses, const pcs_annotate_params* params, pcs_point_xyzrgb* annotated_out,
                        uint8_t* mask_out)
{
  PCS_TRY(ctx_bind(ctx));
  PCS_CHECK_ARG(clouds && params, "NULL argument");
  PCS_CHECK_ARG(n_frames >= 0 && width > 0 && height > 0, "bad frame shape");
  PCS_CHECK_ARG((unsigned long long)width * height < (1ull << 31), "frame too large");
  PCS_CHECK_ARG(params->lower <= params->upper, "colorPassthrough: lower limit greater than upper limit");
  if (n_frames == 0)
    return PCS_OK;

  // Chunked, double-buffered pipeline: H2D on copy stream 0, kernel on the context stream, D2H on copy stream 1.
  const size_t ppf = (size_t)width * height;
  const size_t frame_bytes = ppf * sizeof(pcs_point_xyzrgb);
  size_t chunk_frames = (size_t)(64u << 20) / frame_bytes;
  if (chunk_frames < 1)
    chunk_frames = 1;
  if (chunk_frames > (size_t)n_frames)
    chunk_frames = (size_t)n_frames;
  const size_t n_chunks = ((size_t)n_frames + chunk_frames - 1) / chunk_frames;
  const int n_buf = n_chunks > 1 ? 2 : 1;

  void *d_in[2] = { nullptr, nullptr }, *d_ann[2] = { nullptr, nullptr }, *d_mask[2] = { nullptr, nullptr };
  void* d_pose = nullptr;
  for (int s = 0; s < n_buf; s++)
  {
    void* base = nullptr;
    // one staging slot per buffer set: [in | ann | mask]
    PCS_TRY(ctx_staging(ctx, s, chunk_frames * (2 * frame_bytes + ppf) + 256, &base));
    d_in[s] = base;
    d_ann[s] = (char*)base + chunk_frames * frame_bytes;
    d_mask[s] = (char*)base + 2 * chunk_frames * frame_bytes;
  }
  if (poses)
  {
    PCS_TRY(ctx_staging(ctx, 2, (size_t)n_frames * 16 * sizeof(float), &d_pose));
    PCS_CUDA(cudaMemcpyAsync(d_pose, poses, (size_t)n_frames * 16 * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
    ctx->stats.bytes_h2d += (size_t)n_frames * 64;
  }
  // events live in the context (created once): a per-call create/destroy would be most of a single frame's latency
  cudaEvent_t* ev_in = ctx->pipe_event;
  cudaEvent_t* ev_k = ctx->pipe_event + 2;
  cudaEvent_t* ev_out = ctx->pipe_event + 4;
  int status = PCS_OK;
  cudaError_t first_err = cudaSuccess;  // first failing runtime call of the pipeline (non-sticky errors included)
  const char* first_what = "";
#define PCS_PIPE(call)                                                                                                 \
  do                                                                                                                   \
  {                                                                                                                    \
    cudaError_t e_ = (call);                                                                                           \
    if (e_ != cudaSuccess && first_err == cudaSuccess)                                                                 \
    {                                                                                                                  \
      first_err = e_;                                                                                                  \
      first_what = #call;                                                                                              \
    }                                                                                                                  \
  } while (0)
  cudaStream_t s_in = ctx->copy_stream[0], s_out = ctx->copy_stream[1];
  for (size_t c = 0; c < n_chunks && status == PCS_OK && first_err == cudaSuccess; c++)
  {
    const int s = (int)(c % (size_t)n_buf);
    const size_t f0 = c * chunk_frames;
    const size_t nf = (f0 + chunk_frames <= (size_t)n_frames) ? chunk_frames : (size_t)n_frames - f0;
    if (c >= (size_t)n_buf)
      PCS_PIPE(cudaStreamWaitEvent(s_in, ev_out[s], 0));  // buffer set free again
    PCS_PIPE(cudaMemcpyAsync(d_in[s], clouds + f0 * ppf, nf * frame_bytes, cudaMemcpyHostToDevice, s_in));
    PCS_PIPE(cudaEventRecord(ev_in[s], s_in));
    PCS_PIPE(cudaStreamWaitEvent(ctx->stream, ev_in[s], 0));
    status = annotate_dev(ctx, map, (const pcs_point_xyzrgb*)d_in[s], (int)nf, width, height,
                          d_pose ? (const float*)d_pose + f0 * 16 : nullptr, params,
                          annotated_out ? (pcs_point_xyzrgb*)d_ann[s] : nullptr, mask_out ? (uint8_t*)d_mask[s] : nullptr);
    PCS_PIPE(cudaEventRecord(ev_k[s], ctx->stream));
    PCS_PIPE(cudaStreamWaitEvent(s_out, ev_k[s], 0));
    if (annotated_out)
      PCS_PIPE(cudaMemcpyAsync(annotated_out + f0 * ppf, d_ann[s], nf * frame_bytes, cudaMemcpyDeviceToHost, s_out));
    if (mask_out)
      PCS_PIPE(cudaMemcpyAsync(mask_out + f0 * ppf, d_mask[s], nf * ppf, cudaMemcpyDeviceToHost, s_out));
    PCS_PIPE(cudaEventRecord(ev_out[s], s_out));
    ctx->stats.bytes_h2d += nf * frame_bytes;
    ctx->stats.bytes_d2h += (annotated_out ? nf * frame_bytes : 0) + (mask_out ? nf * ppf : 0);
  }
#undef PCS_PIPE
  cudaError_t e1 = cudaStreamSynchronize(s_in), e2 = cudaStreamSynchronize(ctx->stream), e3 = cudaStreamSynchronize(s_out);
  if (status != PCS_OK)
    return status;
  if (first_err != cudaSuccess)
    return cuda_fail(first_err, first_what, __FILE__, __LINE__);
  PCS_CUDA(e1);
  PCS_CUDA(e2);
  PCS_CUDA(e3);
  return PCS_OK;
}

int pcs_map_integrate_dev(pcs_ctx* ctx, pcs_map* map, const pcs_point_xyzrgb* points_dev, uint64_t n,
                          const float* pose_dev, int lower, int upper, int limit_negative)
{
  PCS_TRY(ctx_bind(ctx));
  PCS_CHECK_ARG(map != nullptr, "map is NULL");
  PCS_CHECK_ARG(lower <= upper, "colorPassthrough: lower limit greater than upper limit");
  if (n == 0)
    return PCS_OK;
  PCS_CHECK_ARG(points_dev != nullptr, "points is NULL");
  // one 256-point group per CTA up to 2^31 - 1 CTAs (the hardware block scheduler streams better than a persistent loop)
  unsigned long long blocks = (n + 255) / 256;
  if (blocks > 0x7fffffffull)
    blocks = 0x7fffffffull;
  integrate_kernel<<<(unsigned int)blocks, 256, 0, ctx->stream>>>((const Point8*)points_dev, n, pose_dev, lower, upper,
                                                                limit_negative ? 1 : 0, map->view);
  count_launch(ctx);
  ctx->stats.points_in += n;
  return check_launch("integrate_kernel");
}

int pcs_map_integrate(pcs_ctx* ctx, pcs_map* map, const pcs_point_xyzrgb* points, uint64_t n, const float* pose, int lower,
                      int upper, int limit_negative)
{
  PCS_TRY(ctx_bind(ctx));
  PCS_CHECK_ARG(map != nullptr, "map is NULL");
  PCS_CHECK_ARG(lower <= upper, "colorPassthrough: lower limit greater than upper limit");
  if (n == 0)
    return PCS_OK;
  PCS_CHECK_ARG(points != nullptr, "points is NULL");
  void *d_pts = nullptr, *d_pose = nullptr;
  PCS_TRY(ctx_staging(ctx, 0, n * sizeof(pcs_point_xyzrgb), &d_pts));
  PCS_CUDA(cudaMemcpyAsync(d_pts, points, n * sizeof(pcs_point_xyzrgb), cudaMemcpyHostToDevice, ctx->stream));
  if (pose)
  {
    PCS_TRY(ctx_staging(ctx, 2, 64, &d_pose));
    PCS_CUDA(cudaMemcpyAsync(d_pose, pose, 64, cudaMemcpyHostToDevice, ctx->stream));
  }
  ctx->stats.bytes_h2d += n * sizeof(pcs_point_xyzrgb);
  PCS_TRY(pcs_map_integrate_dev(ctx, map, (const pcs_point_xyzrgb*)d_pts, n, (const float*)d_pose, lower, upper,
                                limit_negative));
  PCS_CUDA(cudaStreamSynchronize(ctx->stream));
  return PCS_OK;
}

int pcs_cloud_to_image(pcs_ctx* ctx, const pcs_point_xyzrgb* cloud, int width, int height, double* position_image,
                       uint8_t* bgr_image)
{
  PCS_TRY(ctx_bind(ctx));
  PCS_CHECK_ARG(cloud && position_image && bgr_image, "NULL argument");
  PCS_CHECK_ARG(width > 0 && height > 0, "empty cloud");
  PCS_CHECK_ARG(width != 1 || height != 1, "cloud is not organized (1x1), utils.h:103");
  const size_t n = (size_t)width * height;
  void *d_in = nullptr, *d_out = nullptr;
  PCS_TRY(ctx_staging(ctx, 0, n * 32, &d_in));
  PCS_TRY(ctx_staging(ctx, 1, n * 27, &d_out));
  double* d_pos = (double*)d_out;
  uint8_t* d_bgr = (uint8_t*)d_out + n * 24;
  PCS_CUDA(cudaMemcpyAsync(d_in, cloud, n * 32, cudaMemcpyHostToDevice, ctx->stream));
  cloud_to_image_kernel<<<(unsigned int)((n + 255) / 256), 256, 0, ctx->stream>>>((const Point8*)d_in, n, d_pos, d_bgr);
  count_launch(ctx);
  PCS_TRY(check_launch("cloud_to_image_kernel"));
  PCS_CUDA(cudaMemcpyAsync(position_image, d_pos, n * 24, cudaMemcpyDeviceToHost, ctx->stream));
  PCS_CUDA(cudaMemcpyAsync(bgr_image, d_bgr, n * 3, cudaMemcpyDeviceToHost, ctx->stream));
  PCS_CUDA(cudaStreamSynchronize(ctx->stream));
  ctx->stats.bytes_h2d += n * 32;
  ctx->stats.bytes_d2h += n * 27;
  return PCS_OK;
}

int pcs_image_to_cloud(pcs_ctx* ctx, const uint8_t* annotation_bgr, const double* position_image, int width, int height,
                       pcs_point_xyzrgb* cloud_out)
{
  PCS_TRY(ctx_bind(ctx));
  PCS_CHECK_ARG(annotation_bgr && position_image && cloud_out, "NULL argument");
  PCS_CHECK_ARG(width > 0 && height > 0, "empty image");
  const size_t n = (size_t)width * height;
  void *d_in = nullptr, *d_out = nullptr;
  PCS_TRY(ctx_staging(ctx, 0, n * 27, &d_in));
  PCS_TRY(ctx_staging(ctx, 1, n * 32, &d_out));
  double* d_pos = (double*)d_in;
  uint8_t* d_ann = (uint8_t*)d_in + n * 24;
  PCS_CUDA(cudaMemcpyAsync(d_pos, position_image, n * 24, cudaMemcpyHostToDevice, ctx->stream));
  PCS_CUDA(cudaMemcpyAsync(d_ann, annotation_bgr, n * 3, cudaMemcpyHostToDevice, ctx->stream));
  image_to_cloud_kernel<<<(unsigned int)((n + 255) / 256), 256, 0, ctx->stream>>>(d_ann, d_pos, n, (Point8*)d_out);
  count_launch(ctx);
  PCS_TRY(check_launch("image_to_cloud_kernel"));
  PCS_CUDA(cudaMemcpyAsync(cloud_out, d_out, n * 32, cudaMemcpyDeviceToHost, ctx->stream));
  PCS_CUDA(cudaStreamSynchronize(ctx->stream));
  ctx->stats.bytes_h2d += n * 27;
  ctx->stats.bytes_d2h += n * 32;
  return PCS_OK;
}

int pcs_transform_cloud(pcs_ctx* ctx, const float* pose, const pcs_point_xyzrgb* in, uint64_t n, pcs_point_xyzrgb* out)
{
  PCS_TRY(ctx_bind(ctx));
  PCS_CHECK_ARG(pose != nullptr, "pose is NULL");
  if (n == 0)
    return PCS_OK;
  PCS_CHECK_ARG(in && out, "NULL argument");
  void *d_in = nullptr, *d_out = nullptr, *d_pose = nullptr;
  PCS_TRY(ctx_staging(ctx, 0, n * 32, &d_in));
  PCS_TRY(ctx_staging(ctx, 1, n * 32, &d_out));
  PCS_TRY(ctx_staging(ctx, 2, 64, &d_pose));
  PCS_CUDA(cudaMemcpyAsync(d_in, in, n * 32, cudaMemcpyHostToDevice, ctx->stream));
  PCS_CUDA(cudaMemcpyAsync(d_pose, pose, 64, cudaMemcpyHostToDevice, ctx->stream));
  transform_kernel<<<(unsigned int)((n + 255) / 256), 256, 0, ctx->stream>>>((const Point8*)d_in, n, (const float*)d_pose,
                                                                            (Point8*)d_out);
  count_launch(ctx);
  PCS_TRY(check_launch("transform_kernel"));
  PCS_CUDA(cudaMemcpyAsync(out, d_out, n * 32, cudaMemcpyDeviceToHost, ctx->stream));
  PCS_CUDA(cudaStreamSynchronize(ctx->stream));
  ctx->stats.bytes_h2d += n * 32 + 64;
  ctx->stats.bytes_d2h += n * 32;
  return PCS_OK;
}

int pcs_labelled_points(pcs_ctx* ctx, const pcs_point_xyzrgb* clouds, int n_frames, int width, int height,
                        const float* poses, const pcs_annotate_params* params, double resolution, uint64_t cap,
                        pcs_point_xyzrgb* out, uint64_t* n_out)
{
  PCS_TRY(ctx_bind(ctx));
  PCS_CHECK_ARG(params && n_out, "NULL argument");
  PCS_CHECK_ARG(n_frames >= 0 && width > 0 && height > 0, "bad frame shape");
  PCS_CHECK_ARG(params->lower <= params->upper, "colorPassthrough: lower limit greater than upper limit");
  PCS_CHECK_ARG(resolution >= 0.0, "resolution must be >= 0");
  *n_out = 0;
  const uint64_t n = (uint64_t)n_frames * width * height;
  if (n == 0)
    return PCS_OK;
  PCS_CHECK_ARG(clouds != nullptr, "clouds is NULL");
  void *d_in = nullptr, *d_ann = nullptr, *d_out = nullptr, *d_pose = nullptr, *scr = nullptr, *h_total = nullptr;
  PCS_TRY(ctx_staging(ctx, 0, n * 32, &d_in));
  PCS_TRY(ctx_staging(ctx, 1, n * 32, &d_ann));
  PCS_TRY(ctx_staging(ctx, 3, n * 32, &d_out));
  PCS_TRY(ctx_scratch(ctx, compact_scratch_bytes(n), &scr));
  PCS_TRY(ctx_pinned(ctx, 8, &h_total));
  PCS_CUDA(cudaMemcpyAsync(d_in, clouds, n * 32, cudaMemcpyHostToDevice, ctx->stream));
  if (poses)
  {
    PCS_TRY(ctx_staging(ctx, 2, (size_t)n_frames * 64, &d_pose));
    PCS_CUDA(cudaMemcpyAsync(d_pose, poses, (size_t)n_frames * 64, cudaMemcpyHostToDevice, ctx->stream));
  }
  // pass 1: detector + labelling + transform (world-frame annotated cloud); pass 2: ballot compaction of the survivors
  pcs_annotate_params prm = *params;
  prm.world_frame_output = 1;
  PCS_TRY(annotate_dev(ctx, nullptr, (const pcs_point_xyzrgb*)d_in, n_frames, width, height, (const float*)d_pose, &prm,
                       (pcs_point_xyzrgb*)d_ann, nullptr));
  LabelledPred pred{ (const Point8*)d_ann, resolution > 0.0 ? 1.0 / resolution : 0.0, params->lower, params->upper,
                     params->limit_negative ? 1 : 0 };
  PassEmit emit{ (const Point8*)d_ann, (Point8*)d_out };
  const unsigned long long* d_total = nullptr;
  PCS_TRY(compact_run(ctx, pred, emit, n, scr, &d_total));
  PCS_CUDA(cudaMemcpyAsync(h_total, d_total, 8, cudaMemcpyDeviceToHost, ctx->stream));
  PCS_CUDA(cudaStreamSynchronize(ctx->stream));
  const uint64_t m = *(const unsigned long long*)h_total;
  *n_out = m;
  ctx->stats.bytes_h2d += n * 32;
  if (!out)
    return PCS_OK;  // count query
  if (m > cap)
  {
    set_error("output holds %llu points, %llu labelled points found", (unsigned long long)cap, (unsigned long long)m);
    return PCS_ERR_CAPACITY;
  }
  if (m)
    PCS_CUDA(cudaMemcpyAsync(out, d_out, m * 32, cudaMemcpyDeviceToHost, ctx->stream));
  PCS_CUDA(cudaStreamSynchronize(ctx->stream));
  ctx->stats.bytes_d2h += m * 32 + 8;
  return PCS_OK;
}

int pcs_color_passthrough(pcs_ctx* ctx, const pcs_point_xyzrgb* in, uint64_t n, int lower, int upper, int limit_negative,
                          pcs_point_xyzrgb* out, uint64_t* n_out)
{
  PCS_TRY(ctx_bind(ctx));
  PCS_CHECK_ARG(n_out != nullptr, "n_out is NULL");
  PCS_CHECK_ARG(lower <= upper, "colorPassthrough: lower limit greater than upper limit");
  *n_out = 0;
  if (n == 0)
    return PCS_OK;
  PCS_CHECK_ARG(in && out, "NULL argument");
  void *d_in = nullptr, *d_out = nullptr, *scr = nullptr, *h_total = nullptr;
  PCS_TRY(ctx_staging(ctx, 0, n * 32, &d_in));
  PCS_TRY(ctx_staging(ctx, 1, n * 32, &d_out));
  PCS_TRY(ctx_scratch(ctx, compact_scratch_bytes(n), &scr));
  PCS_TRY(ctx_pinned(ctx, 8, &h_total));
  PCS_CUDA(cudaMemcpyAsync(d_in, in, n * 32, cudaMemcpyHostToDevice, ctx->stream));
  PassPred pred{ (const Point8*)d_in, lower, upper, limit_negative ? 1 : 0 };
  PassEmit emit{ (const Point8*)d_in, (Point8*)d_out };
  const unsigned long long* d_total = nullptr;
  PCS_TRY(compact_run(ctx, pred, emit, n, scr, &d_total));
  PCS_CUDA(cudaMemcpyAsync(h_total, d_total, 8, cudaMemcpyDeviceToHost, ctx->stream));
  PCS_CUDA(cudaStreamSynchronize(ctx->stream));
  const uint64_t m = *(const unsigned long long*)h_total;
  if (m)
    PCS_CUDA(cudaMemcpyAsync(out, d_out, m * 32, cudaMemcpyDeviceToHost, ctx->stream));
  PCS_CUDA(cudaStreamSynchronize(ctx->stream));
  *n_out = m;
  ctx->stats.bytes_h2d += n * 32;
  ctx->stats.bytes_d2h += m * 32 + 8;
  return PCS_OK;
}

}  // extern "C"
