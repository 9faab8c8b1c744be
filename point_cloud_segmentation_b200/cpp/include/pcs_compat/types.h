// pcs_compat/types.h -- layout-compatible stand-ins for the third-party types in the reference's public signatures.
//
// The reference's headers speak pcl::PointCloud<pcl::PointXYZRGB>, cv::Mat, tesseract_geometry::Mesh / Octree
// (pcs_detection/include/pcs_detection/utils.h:29-33, pcs_scan_integration/include/pcs_scan_integration/
// octomap_mesh_masking.h:30-34).  None of those libraries exists in this build environment, so the adapters are
// written against the minimal types below, which copy only the members the hot path touches and keep the memory
// layout that crosses the C ABI (sizeof(PointXYZRGB) == 32, Mat rows with a byte step, faces as [3,i0,i1,i2]).
// Where PCL / OpenCV / tesseract are installed, alias these names to the real types instead.
#pragma once

#include <array>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "pcs_b200.h"

namespace pcs_compat
{
// ---- pcl ---------------------------------------------------------------------------------------------------------
struct alignas(16) PointXYZRGB
{
  float x, y, z, data3;
  union
  {
    struct
    {
      uint8_t b, g, r, a;
    };
    float rgb;
    uint32_t rgba;
  };
  uint32_t pad[3];
  // pcl::PointXYZRGB(): x = y = z = 0, data[3] = 1, r = g = b = 0, a = 255
  PointXYZRGB() : x(0.f), y(0.f), z(0.f), data3(1.f), rgba(0xff000000u), pad{ 0, 0, 0 } {}
};
static_assert(sizeof(PointXYZRGB) == sizeof(pcs_point_xyzrgb) && sizeof(PointXYZRGB) == 32, "PointXYZRGB layout");

struct PCLHeader
{
  uint32_t seq = 0;
  uint64_t stamp = 0;
  std::string frame_id;
};

template <typename PointT>
struct PointCloud
{
  typedef std::shared_ptr<PointCloud<PointT>> Ptr;
  typedef std::shared_ptr<const PointCloud<PointT>> ConstPtr;
  PCLHeader header;
  std::vector<PointT> points;
  uint32_t width = 0;
  uint32_t height = 0;
  bool is_dense = true;
  std::size_t size() const { return points.size(); }
  void resize(std::size_t n)
  {
    points.resize(n);
    if (width * height != n)
    {
      width = static_cast<uint32_t>(n);
      height = 1;
    }
  }
};

// ---- OpenCV ----------------------------------------------------------------------------------------------------------
enum MatType
{
  CV_8UC1 = 0,
  CV_8UC3 = 16,
  CV_64FC3 = 22
};

inline std::size_t elemSize(int type) { return type == CV_8UC1 ? 1 : (type == CV_8UC3 ? 3 : 24); }

// Dense 2D array with shared, reference-counted storage (cv::Mat's copy semantics: copies share data, clone() is deep).
struct Mat
{
  int rows = 0, cols = 0;
  int mat_type = CV_8UC1;
  std::size_t step = 0;
  uint8_t* data = nullptr;
  std::shared_ptr<std::vector<uint8_t>> storage;

  Mat() = default;
  Mat(int r, int c, int t) { create(r, c, t); }
  void create(int r, int c, int t)
  {
    rows = r;
    cols = c;
    mat_type = t;
    step = static_cast<std::size_t>(c) * elemSize(t);
    storage = std::make_shared<std::vector<uint8_t>>(step * static_cast<std::size_t>(r));
    data = storage->data();
  }
  int type() const { return mat_type; }
  bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
  Mat clone() const
  {
    Mat m(rows, cols, mat_type);
    if (!empty())
      std::memcpy(m.data, data, step * static_cast<std::size_t>(rows));
    return m;
  }
  // cv::Mat::at<T>(row, col) with col counted in elements of T (so position_image.at<double>(y, x * 3 + c) works)
  template <typename T>
  T& at(int r, int c)
  {
    return *reinterpret_cast<T*>(data + step * static_cast<std::size_t>(r) + sizeof(T) * static_cast<std::size_t>(c));
  }
  template <typename T>
  const T& at(int r, int c) const
  {
    return *reinterpret_cast<const T*>(data + step * static_cast<std::size_t>(r) + sizeof(T) * static_cast<std::size_t>(c));
  }
};

// ---- tesseract_geometry ------------------------------------------------------------------------------------------------
typedef std::array<double, 3> Vector3d;
typedef std::array<int, 3> Vector3i;
typedef std::vector<Vector3d> VectorVector3d;
typedef std::vector<int> VectorXi;

class Mesh
{
public:
  typedef std::shared_ptr<Mesh> Ptr;
  typedef std::shared_ptr<const Mesh> ConstPtr;
  Mesh(std::shared_ptr<const VectorVector3d> vertices, std::shared_ptr<const VectorXi> triangles)
    : vertices_(std::move(vertices)), triangles_(std::move(triangles))
  {
    triangle_count_ = static_cast<int>(triangles_->size() / 4);
  }
  const std::shared_ptr<const VectorVector3d>& getVertices() const { return vertices_; }
  const std::shared_ptr<const VectorXi>& getTriangles() const { return triangles_; }  // [3, i0, i1, i2] per face
  int getVerticeCount() const { return static_cast<int>(vertices_->size()); }
  int getTriangleCount() const { return triangle_count_; }

private:
  std::shared_ptr<const VectorVector3d> vertices_;
  std::shared_ptr<const VectorXi> triangles_;
  int triangle_count_;
};

// One process-wide context on the device named by $PCS_B200_DEVICE (default 0): the reference's functions are free
// functions without a handle argument, so the adapters need an implicit one.
inline pcs_ctx* default_ctx()
{
  struct Holder
  {
    pcs_ctx* ctx = nullptr;
    Holder()
    {
      const char* dev = std::getenv("PCS_B200_DEVICE");
      if (pcs_ctx_create(dev ? std::atoi(dev) : 0, nullptr, &ctx) != PCS_OK)
        throw std::runtime_error(std::string("libpcs_b200: ") + pcs_last_error());
    }
    ~Holder() { pcs_ctx_destroy(ctx); }
  };
  static Holder h;
  return h.ctx;
}

// console_bridge stand-in
inline void logError(const char* what) { std::fprintf(stderr, "Error:   %s\n", what); }
}  // namespace pcs_compat
