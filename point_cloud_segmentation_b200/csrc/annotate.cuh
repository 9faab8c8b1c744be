// annotate.cuh -- per-point body shared by the fused annotate(+integrate) kernels (annotate.cu: plain HSV detector;
// annotate_open.cu: HSV detector followed by the 5x5 opening).  See annotate.cu for the reference call chain it replaces.
#pragma once

#include "common.cuh"
#include "hsv.cuh"

namespace pcs
{
struct AnnotateArgs
{
  const Point8* cloud;
  Point8* ann;
  uint8_t* mask;
  const float* poses;      // n_frames x 16 row-major, or NULL
  unsigned int pts_per_frame;
  unsigned int tiles_per_frame;
  unsigned int n_tiles;
  unsigned int width, height;  // frame shape (the opening kernel works on the organized grid)
  HsvBounds hsv;
  const HsvDecisionTables* hsv_tables;  // device tables for `hsv` (annotate_open_kernel), else NULL
  int inverted;
  int pass_black, pass_white;  // colorPassthrough of the two annotation colours
  int world_out;
  MapView map;
};

struct Pose
{
  float m[12];
};

// Outputs of one point once its 32 bytes have arrived and the detector's verdict `det` is known: annotated point and
// mask byte are stored; the voxel key, its validity and its label class are returned for the map update.
template <bool HAS_MAP, bool WRITE_ANN, bool WRITE_MASK>
__device__ __forceinline__ void emit_point(const AnnotateArgs& a, const Pose& P, const bool has_pose, const Point8& p,
                                           const unsigned long long idx, const bool in_frame, const bool det, bool& valid,
                                           unsigned long long& key, bool& cls1)
{
  const float x = p.x, y = p.y, z = p.z;
  float wx = x, wy = y, wz = z;
  if (has_pose && (HAS_MAP || WRITE_ANN))
  {
    // non-finite inputs come out non-finite in every component (NaN * 0 = NaN), so no test is needed before
    wx = xform_row(P.m[0], P.m[1], P.m[2], P.m[3], x, y, z);
    wy = xform_row(P.m[4], P.m[5], P.m[6], P.m[7], x, y, z);
    wz = xform_row(P.m[8], P.m[9], P.m[10], P.m[11], x, y, z);
  }
  if (in_frame)
  {
    if (WRITE_MASK)
      a.mask[idx] = det ? 255 : 0;
    if (WRITE_ANN)
    {
      Point8 o;
      // pcl::transformPointCloud leaves non-finite points untouched
      const bool w = a.world_out != 0 && finite3(x, y, z);
      o.x = w ? wx : x;
      o.y = w ? wy : y;
      o.z = w ? wz : z;
      o.w = 1.0f;                       // pcl::PointXYZRGB() sets data[3] = 1
      o.rgba = det ? 0x00ffffffu : 0u;  // r = g = b = mask, alpha 0 (utils.h:161-162)
      o.p0 = o.p1 = o.p2 = 0u;
      st_point_stream(a.ann + idx, o);
    }
  }
  valid = false;
  key = 0ull;
  cls1 = false;
  if (HAS_MAP)
  {
    unsigned int kx, ky, kz;
    valid = in_frame && finite3(wx, wy, wz);
    valid = coord_to_key(a.map.inv_res, wx, kx) && valid;
    valid = coord_to_key(a.map.inv_res, wy, ky) && valid;
    valid = coord_to_key(a.map.inv_res, wz, kz) && valid;
    key = ((unsigned long long)kz << 32) | ((unsigned long long)ky << 16) | kx;
    cls1 = det ? (a.pass_white != 0) : (a.pass_black != 0);
  }
}

template <bool HAS_MAP, bool WRITE_ANN, bool WRITE_MASK>
__device__ __forceinline__ void finish_point(const AnnotateArgs& a, const Pose& P, const bool has_pose, const Point8& p,
                                             const unsigned long long idx, const bool in_frame, const bool det)
{
  bool valid, cls1;
  unsigned long long key;
  emit_point<HAS_MAP, WRITE_ANN, WRITE_MASK>(a, P, has_pose, p, idx, in_frame, det, valid, key, cls1);
  if (HAS_MAP)
    map_add_warp(a.map, valid, key, cls1);
}

// plain detector: threshold the point's own colour
template <bool HAS_MAP, bool WRITE_ANN, bool WRITE_MASK>
__device__ __forceinline__ void annotate_point(const AnnotateArgs& a, const int* __restrict__ s_lut, const Pose& P,
                                               const bool has_pose, const Point8& p, const unsigned long long idx,
                                               const bool in_frame)
{
  const int b = p.rgba & 0xff, g = (p.rgba >> 8) & 0xff, r = (p.rgba >> 16) & 0xff;
  const bool det = hsv_in_range(s_lut, b, g, r, a.hsv) != (a.inverted != 0);
  finish_point<HAS_MAP, WRITE_ANN, WRITE_MASK>(a, P, has_pose, p, idx, in_frame, det);
}

__device__ __forceinline__ void load_pose(const float* __restrict__ poses, unsigned int frame, Pose& P)
{
  const float4* src = reinterpret_cast<const float4*>(poses + (size_t)frame * 16);
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

// fills the argument block both kernels share; returns PCS_OK or an error status
int fill_annotate_args(pcs_map* map, const Point8* clouds, unsigned int n_frames, unsigned int width, unsigned int height,
                       const float* poses, const pcs_annotate_params* prm, Point8* ann, uint8_t* mask, AnnotateArgs& a);
int launch_annotate_open(pcs_ctx* ctx, const AnnotateArgs& a, unsigned int n_frames, bool has_map, bool write_ann,
                         bool write_mask);
}  // namespace pcs
