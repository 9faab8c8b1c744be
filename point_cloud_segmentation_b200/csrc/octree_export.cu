// octree_export.cu -- OcTree-compatible views of the voxel map.
//
//   pcs_map_write_bt       the occupied voxels (count1 >= 1) as an octomap binary tree (.bt), the file format the
//                          reference's test loads with octomap::OcTree(path)
//                          (pcs_scan_integration/test/octomap_mesh_masking_unit.cpp:140-141).  Same bytes as
//                          octomap's OcTree::writeBinary (toMaxLikelihood + prune + 2 bits per child) for the same
//                          occupied set: complete octets of occupied leaves are collapsed bottom-up (depth 15 .. 1,
//                          never into the root), children are written in child-index order (bit0 x, bit1 y, bit2 z).
//                          One documented difference: octomap's prune() stops at the first depth where nothing was
//                          collapsed, so a tree whose depth-15 octets were all merged during insertion can keep an
//                          uncollapsed complete octet one level up; this writer always collapses fully.
//   pcs_map_leaf_centers   one PointXYZRGB per voxel at the voxel centre, white where count1 >= 1 else black -- the
//                          "octomap as a colourised centre cloud" that the mesh-mask action ingests
//                          (pcs_scan_integration/README.md:11-13, pcs_ros/src/octomap_mesh_mask_node.cpp:47-68).
//                          setOctree(centres, resolution, 255, 255) rebuilds the occupied set exactly.
//
// The device part is the Morton-ordered export (voxel_map.cu); tree assembly is sequential host work on <= 10^6 keys.
#include <stdio.h>
#include <string.h>

#include <string>
#include <vector>

#include "common.cuh"

namespace pcs
{
static inline unsigned long long spread16_host(unsigned long long x)
{
  x &= 0xffffull;
  x = (x | (x << 16)) & 0x000000ff0000ffull;
  x = (x | (x << 8)) & 0x00f00f00f00f00full;
  x = (x | (x << 4)) & 0x0c30c30c30c30c3ull;
  x = (x | (x << 2)) & 0x249249249249249ull;
  return x;
}

static inline unsigned long long morton48_host(unsigned long long key48)
{
  return spread16_host(key48) | (spread16_host(key48 >> 16) << 1) | (spread16_host(key48 >> 32) << 2);
}

struct Leaf
{
  unsigned long long prefix;  // Morton code >> 3 * (16 - depth)
  int depth;                  // 16 = finest voxel
};

// leaves arrive in DFS (Morton) order; collapse complete octets of equal-depth leaves, deepest level first
static void prune_leaves(std::vector<Leaf>& leaves)
{
  for (int d = 16; d >= 2; --d)
  {
    std::vector<Leaf> out;
    out.reserve(leaves.size());
    size_t i = 0;
    while (i < leaves.size())
    {
      if (leaves[i].depth == d && (leaves[i].prefix & 7ull) == 0ull && i + 8 <= leaves.size())
      {
        bool full = true;
        for (int c = 0; c < 8 && full; c++)
          full = leaves[i + c].depth == d && leaves[i + c].prefix == leaves[i].prefix + (unsigned long long)c;
        if (full)
        {
          out.push_back(Leaf{ leaves[i].prefix >> 3, d - 1 });
          i += 8;
          continue;
        }
      }
      out.push_back(leaves[i++]);
    }
    leaves.swap(out);
  }
}

struct BtWriter
{
  const std::vector<Leaf>& leaves;
  std::vector<uint8_t> bytes;
  size_t n_nodes = 0;

  // leaves[lo, hi) all lie under the node (prefix, depth), which is an inner node
  void write_inner(unsigned long long prefix, int depth, size_t lo, size_t hi)
  {
    n_nodes++;
    size_t begin[9];
    size_t p = lo;
    for (int c = 0; c < 8; c++)
    {
      begin[c] = p;
      const unsigned long long child = (prefix << 3) | (unsigned long long)c;
      while (p < hi && (leaves[p].prefix >> (3 * (leaves[p].depth - (depth + 1)))) == child)
        p++;
    }
    begin[8] = p;
    unsigned bits = 0;
    for (int c = 0; c < 8; c++)
    {
      const size_t n = begin[c + 1] - begin[c];
      if (n == 0)
        continue;
      const bool is_leaf = n == 1 && leaves[begin[c]].depth == depth + 1;
      bits |= (is_leaf ? 2u : 3u) << (2 * c);  // bit pair (child[2c], child[2c+1]): 0/1 occupied leaf, 1/1 inner node
      if (is_leaf)
        n_nodes++;
    }
    bytes.push_back((uint8_t)(bits & 0xff));
    bytes.push_back((uint8_t)(bits >> 8));
    for (int c = 0; c < 8; c++)
      if (((bits >> (2 * c)) & 3u) == 3u)
        write_inner((prefix << 3) | (unsigned long long)c, depth + 1, begin[c], begin[c + 1]);
  }
};

__global__ void leaf_centers_kernel(const unsigned long long* __restrict__ key48, const unsigned int* __restrict__ c1,
                                    unsigned int n, double res, Point8* __restrict__ out)
{
  const unsigned int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n)
    return;
  const unsigned long long k = key48[i];
  Point8 p;
  // octomap keyToCoord: (key - 32768 + 0.5) * resolution, narrowed to float as octomap::point3d does
  p.x = (float)__dmul_rn((double)((int)(k & 0xffff) - 32768) + 0.5, res);
  p.y = (float)__dmul_rn((double)((int)((k >> 16) & 0xffff) - 32768) + 0.5, res);
  p.z = (float)__dmul_rn((double)((int)((k >> 32) & 0xffff) - 32768) + 0.5, res);
  p.w = 1.0f;
  p.rgba = c1[i] ? 0xffffffffu : 0xff000000u;  // opaque white = labelled, opaque black = seen but not labelled
  p.p0 = p.p1 = p.p2 = 0u;
  st_point_stream(out + i, p);
}
}  // namespace pcs

using namespace pcs;

extern "C" {

int pcs_map_write_bt(pcs_ctx* ctx, pcs_map* map, uint8_t* out, uint64_t cap, uint64_t* n_bytes)
{
  PCS_TRY(ctx_bind(ctx));
  PCS_CHECK_ARG(map && n_bytes, "NULL argument");
  uint64_t n = 0;
  PCS_TRY(pcs_map_export(ctx, map, 0, nullptr, nullptr, nullptr, nullptr, &n));
  std::vector<uint64_t> keys(n);
  std::vector<uint32_t> c1(n);
  if (n)
    PCS_TRY(pcs_map_export(ctx, map, n, keys.data(), nullptr, c1.data(), nullptr, &n));
  std::vector<Leaf> leaves;
  leaves.reserve(n);
  for (uint64_t i = 0; i < n; i++)
    if (c1[i])
      leaves.push_back(Leaf{ morton48_host(keys[i]), 16 });  // export order is Morton order already
  prune_leaves(leaves);
  BtWriter w{ leaves, {}, 0 };
  if (!leaves.empty())
    w.write_inner(0ull, 0, 0, leaves.size());
  char hdr[256];
  const int hl = snprintf(hdr, sizeof(hdr),
                          "# Octomap OcTree binary file\n# (feel free to add / change comments, but leave the first line as it "
                          "is!)\n#\nid OcTree\nsize %zu\nres %g\ndata\n",
                          w.n_nodes, map->resolution);
  const uint64_t total = (uint64_t)hl + w.bytes.size();
  *n_bytes = total;
  if (!out)
    return PCS_OK;  // size query
  if (cap < total)
  {
    set_error("buffer holds %llu bytes, the .bt needs %llu", (unsigned long long)cap, (unsigned long long)total);
    return PCS_ERR_CAPACITY;
  }
  memcpy(out, hdr, (size_t)hl);
  if (!w.bytes.empty())
    memcpy(out + hl, w.bytes.data(), w.bytes.size());
  return PCS_OK;
}

int pcs_map_leaf_centers(pcs_ctx* ctx, pcs_map* map, uint64_t cap, pcs_point_xyzrgb* points, uint64_t* n_out)
{
  PCS_TRY(ctx_bind(ctx));
  PCS_CHECK_ARG(map && n_out, "NULL argument");
  uint64_t n = 0;
  PCS_TRY(pcs_map_export(ctx, map, 0, nullptr, nullptr, nullptr, nullptr, &n));
  *n_out = n;
  if (!points || n == 0)
    return PCS_OK;
  if (cap < n)
  {
    set_error("buffer holds %llu points, the map has %llu voxels", (unsigned long long)cap, (unsigned long long)n);
    return PCS_ERR_CAPACITY;
  }
  // Morton-ordered keys and counts, then one kernel builds the points on the device
  std::vector<uint64_t> keys(n);
  std::vector<uint32_t> c1(n);
  PCS_TRY(pcs_map_export(ctx, map, n, keys.data(), nullptr, c1.data(), nullptr, &n));
  void* d = nullptr;
  PCS_TRY(ctx_staging(ctx, 0, n * (8 + 4 + 32) + 1024, &d));
  unsigned long long* d_k = (unsigned long long*)d;
  Point8* d_p = (Point8*)((char*)d + ((n * 8 + 255) & ~(size_t)255));
  unsigned int* d_c = (unsigned int*)((char*)d_p + n * 32);
  PCS_CUDA(cudaMemcpyAsync(d_k, keys.data(), n * 8, cudaMemcpyHostToDevice, ctx->stream));
  PCS_CUDA(cudaMemcpyAsync(d_c, c1.data(), n * 4, cudaMemcpyHostToDevice, ctx->stream));
  leaf_centers_kernel<<<(unsigned int)((n + 255) / 256), 256, 0, ctx->stream>>>(d_k, d_c, (unsigned int)n, map->resolution, d_p);
  count_launch(ctx);
  PCS_TRY(check_launch("leaf_centers_kernel"));
  PCS_CUDA(cudaMemcpyAsync(points, d_p, n * 32, cudaMemcpyDeviceToHost, ctx->stream));
  PCS_CUDA(cudaStreamSynchronize(ctx->stream));
  ctx->stats.bytes_d2h += n * 32;
  return PCS_OK;
}

}  // extern "C"
