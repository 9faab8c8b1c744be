// pcs_scan_integration/octomap_mesh_masking.h -- drop-in for
// pcs_scan_integration/include/pcs_scan_integration/octomap_mesh_masking.h:50-150 and
// pcs_scan_integration/src/pcs_scan_integration/octomap_mesh_masking.cpp:52-292.
//
// colorPassthrough and OctomapMeshMask keep their names, arguments, defaults and return conventions.  The octree is
// a device voxel map (pcs_map) wrapped in an Octree stand-in; maskMesh runs pcs_mesh_mask.  RETURN_INSIDE lists each
// contacted triangle once (the reference lists it once per Bullet contact, .cpp:196-216).
#ifndef PCS_SCAN_INTEGRATION_OCTOMAP_MESH_MASKING_H
#define PCS_SCAN_INTEGRATION_OCTOMAP_MESH_MASKING_H

#include "pcs_compat/mesh_io.h"
#include "pcs_compat/types.h"

namespace pcs_scan_integration
{
using pcs_compat::Mesh;
using pcs_compat::PointCloud;
using pcs_compat::PointXYZRGB;

/** @brief Stand-in for tesseract_geometry::Octree: owns the device voxel map that plays the octree's role. */
class Octree
{
public:
  typedef std::shared_ptr<Octree> Ptr;
  typedef std::shared_ptr<const Octree> ConstPtr;
  enum class SubType
  {
    BOX
  };

  /** @brief tesseract_geometry::Octree(cloud, resolution, BOX, prune): one hit per finite in-range point. */
  Octree(const PointCloud<PointXYZRGB>& cloud, double resolution, SubType = SubType::BOX, bool /*prune*/ = true)
    : resolution_(resolution)
  {
    create(2 * cloud.points.size() + 1024);
    if (pcs_map_integrate(pcs_compat::default_ctx(), map_, reinterpret_cast<const pcs_point_xyzrgb*>(cloud.points.data()),
                          cloud.points.size(), nullptr, 0, 255, 0) != PCS_OK)
      throw std::runtime_error(pcs_last_error());
  }

  /** @brief From an octomap binary tree file (octomap::OcTree(path) + Octree(ot, BOX) in the reference's test). */
  explicit Octree(const std::string& bt_path)
  {
    pcs_compat::BtTree bt;
    if (!pcs_compat::readBt(bt_path, bt))
      throw std::runtime_error("cannot read octomap .bt file " + bt_path);
    resolution_ = bt.resolution;
    std::vector<uint64_t> keys;
    for (const pcs_compat::BtLeaf& l : bt.occupied)  // a pruned leaf is the union of its finest-level voxels
    {
      const uint32_t s = 1u << (16 - l.depth);
      for (uint32_t dz = 0; dz < s; dz++)
        for (uint32_t dy = 0; dy < s; dy++)
          for (uint32_t dx = 0; dx < s; dx++)
            keys.push_back((static_cast<uint64_t>(l.kz + dz) << 32) | (static_cast<uint64_t>(l.ky + dy) << 16) | (l.kx + dx));
    }
    create(keys.size() + 1024);
    std::vector<uint32_t> ones(keys.size(), 1u);
    if (pcs_map_merge(pcs_compat::default_ctx(), map_, keys.data(), nullptr, ones.data(), keys.size()) != PCS_OK)
      throw std::runtime_error(pcs_last_error());
  }

  ~Octree()
  {
    if (map_)
      pcs_map_destroy(pcs_compat::default_ctx(), map_);
  }
  Octree(const Octree&) = delete;
  Octree& operator=(const Octree&) = delete;

  pcs_map* map() const { return map_; }
  double resolution() const { return resolution_; }
  /** @brief number of occupied finest-level voxels */
  uint64_t occupiedVoxels() const
  {
    uint64_t nv = 0, no = 0;
    if (pcs_map_size(pcs_compat::default_ctx(), map_, &nv, &no) != PCS_OK)
      throw std::runtime_error(pcs_last_error());
    return no;
  }

  /** @brief octomap::OcTree::writeBinary: the occupied voxels as a .bt file. */
  bool writeBinary(const std::string& path) const
  {
    uint64_t n = 0;
    if (pcs_map_write_bt(pcs_compat::default_ctx(), map_, nullptr, 0, &n) != PCS_OK)
      return false;
    std::vector<uint8_t> bytes(n);
    if (pcs_map_write_bt(pcs_compat::default_ctx(), map_, bytes.data(), n, &n) != PCS_OK)
      return false;
    std::ofstream f(path, std::ios::binary);
    f.write(reinterpret_cast<const char*>(bytes.data()), static_cast<std::streamsize>(bytes.size()));
    return !f.fail();
  }

private:
  void create(std::size_t capacity)
  {
    if (pcs_map_create(pcs_compat::default_ctx(), resolution_, capacity, &map_) != PCS_OK)
      throw std::runtime_error(pcs_last_error());
  }
  pcs_map* map_ = nullptr;
  double resolution_ = 0.0;
};

/**
 * @brief Colour filter for XYZRGB clouds: a point passes when r, g and b are all within [lower, upper]
 * (limit_negative: all outside or on the bounds).  Order-preserving.  Reference .cpp:52-96.
 */
inline PointCloud<PointXYZRGB>::Ptr colorPassthrough(const PointCloud<PointXYZRGB>::ConstPtr input_cloud, const int& lower_limit,
                                                     const int& upper_limit, const bool& limit_negative = false)
{
  PointCloud<PointXYZRGB>::Ptr cloud_filtered(new PointCloud<PointXYZRGB>());
  cloud_filtered->header = input_cloud->header;
  cloud_filtered->width = input_cloud->width;
  cloud_filtered->height = input_cloud->height;
  cloud_filtered->is_dense = input_cloud->is_dense;
  cloud_filtered->points.resize(input_cloud->points.size());
  if (lower_limit > upper_limit)
  {
    pcs_compat::logError("colorPassthrough: Lower limit greater than upper limit");
    throw std::invalid_argument("colorPassthrough: Lower limit greater than upper limit");  // reference: assert(false)
  }
  uint64_t n = 0;
  const int rc = pcs_color_passthrough(pcs_compat::default_ctx(), reinterpret_cast<const pcs_point_xyzrgb*>(input_cloud->points.data()),
                                       input_cloud->points.size(), lower_limit, upper_limit, limit_negative ? 1 : 0,
                                       reinterpret_cast<pcs_point_xyzrgb*>(cloud_filtered->points.data()), &n);
  if (rc != PCS_OK)
    throw std::runtime_error(pcs_last_error());
  cloud_filtered->resize(n);
  return cloud_filtered;
}

class OctomapMeshMask
{
public:
  /** RETURN_INSIDE / RETURN_OUTSIDE / RETURN_COLORIZED, in the reference's order (= the action's 0 / 1 / 2). */
  enum class MaskType
  {
    RETURN_INSIDE,
    RETURN_OUTSIDE,
    RETURN_COLORIZED
  };

  OctomapMeshMask() {}

  /** @brief Sets the input mesh from an absolute file path (binary STL or ASCII PLY). */
  void setInputMesh(std::string& filepath)
  {
    Mesh::Ptr mesh = pcs_compat::createMeshFromPath(filepath);
    if (!mesh)
      throw std::runtime_error("cannot load mesh " + filepath);
    setInputMesh(mesh);
  }
  inline void setInputMesh(Mesh::ConstPtr input_mesh) { input_mesh_ = input_mesh; }
  inline Mesh::ConstPtr getInputMesh() { return input_mesh_; }

  /** @brief Builds the octree from the points of `point_cloud` that pass the colour filter (reference .cpp:105-116). */
  void setOctree(const PointCloud<PointXYZRGB>::ConstPtr point_cloud, const double resolution, const int& lower_limit = 0,
                 const int& upper_limit = 255, const bool& limit_negative = false)
  {
    PointCloud<PointXYZRGB>::Ptr cloud_filtered = colorPassthrough(point_cloud, lower_limit, upper_limit, limit_negative);
    octree_ = std::make_shared<Octree>(*cloud_filtered, resolution, Octree::SubType::BOX, true);
  }
  void setOctree(const Octree::Ptr& octree) { octree_ = octree; }
  inline Octree::ConstPtr getOctree() { return octree_; }

  /** @brief Applies the octree mask to the mesh.  @return true if successful (reference .cpp:118-283). */
  bool maskMesh(const MaskType& mask_type)
  {
    if (!octree_ || !input_mesh_)
    {
      pcs_compat::logError("maskMesh: octree or input mesh not set");
      return false;
    }
    const pcs_compat::VectorVector3d& verts = *input_mesh_->getVertices();
    const pcs_compat::VectorXi& faces = *input_mesh_->getTriangles();
    const uint64_t nv = verts.size(), nt = faces.size() / 4;
    const double* vp = nv ? verts[0].data() : nullptr;
    uint64_t out_nv = 0, out_nt = 0;
    pcs_ctx* ctx = pcs_compat::default_ctx();
    int rc = pcs_mesh_mask(ctx, octree_->map(), vp, nv, faces.data(), nt, static_cast<int>(mask_type), 0, nullptr, nullptr,
                           nullptr, nullptr, &out_nv, &out_nt);
    if (rc != PCS_OK)
    {
      pcs_compat::logError(pcs_last_error());
      return false;
    }
    auto out_verts = std::make_shared<pcs_compat::VectorVector3d>(out_nv);
    auto out_faces = std::make_shared<pcs_compat::VectorXi>(out_nt * 4);
    std::vector<uint8_t> rgb(out_nv * 3);
    rc = pcs_mesh_mask(ctx, octree_->map(), vp, nv, faces.data(), nt, static_cast<int>(mask_type), out_nt ? out_nt : 1,
                       out_nv ? (*out_verts)[0].data() : nullptr, out_faces->data(), rgb.data(), nullptr, &out_nv, &out_nt);
    if (rc != PCS_OK)
    {
      pcs_compat::logError(pcs_last_error());
      return false;
    }
    mesh_vertices_color_.resize(out_nv);
    for (uint64_t i = 0; i < out_nv; i++)
      mesh_vertices_color_[i] = pcs_compat::Vector3i{ rgb[3 * i], rgb[3 * i + 1], rgb[3 * i + 2] };
    masked_mesh_ = std::make_shared<Mesh>(out_verts, out_faces);
    return true;
  }

  inline Mesh::Ptr getMaskedMesh() { return masked_mesh_; }

  /** @brief Saves the last masked mesh as an ASCII PLY with vertex colours (reference .cpp:285-292). */
  bool saveMaskedMesh(std::string& filepath)
  {
    if (!masked_mesh_)
      return false;
    return pcs_compat::writeSimplePlyFile(filepath, *masked_mesh_->getVertices(), mesh_vertices_color_,
                                          *masked_mesh_->getTriangles(), masked_mesh_->getTriangleCount());
  }

protected:
  Mesh::ConstPtr input_mesh_;
  Mesh::Ptr masked_mesh_;
  std::vector<pcs_compat::Vector3i> mesh_vertices_color_;
  Octree::Ptr octree_;
};
}  // namespace pcs_scan_integration
#endif
