// pcs_compat/mesh_io.h -- file I/O at the boundary of the scan-integration adapter (host side, not on the GPU path).
//
// Stands in for what OctomapMeshMask reaches through tesseract / assimp / octomap in the reference:
//   createMeshFromPath(path, scale 1, triangulate)  pcs_scan_integration/src/pcs_scan_integration/octomap_mesh_masking.cpp:98-103
//       (assimp Triangulate | JoinIdenticalVertices: identical vertices are welded, first occurrence keeps its index;
//        pins 6273 / 8 vertices in pcs_scan_integration/test/octomap_mesh_masking_unit.cpp:76-77,85-86)
//   writeSimplePlyFile                              octomap_mesh_masking.cpp:285-292 (ASCII PLY, uchar vertex colours)
//   octomap::OcTree(path) on a .bt file             pcs_scan_integration/test/octomap_mesh_masking_unit.cpp:140-141
#pragma once

#include <cstdio>
#include <fstream>
#include <iomanip>
#include <iterator>
#include <limits>
#include <map>
#include <sstream>
#include <tuple>

#include "pcs_compat/types.h"

namespace pcs_compat
{
inline bool readFile(const std::string& path, std::vector<uint8_t>& bytes)
{
  std::ifstream f(path, std::ios::binary);
  if (!f)
    return false;
  bytes.assign(std::istreambuf_iterator<char>(f), std::istreambuf_iterator<char>());
  return true;
}

class VertexWelder
{
public:
  int add(float x, float y, float z, VectorVector3d& verts)
  {
    auto norm = [](float f) { return f == 0.0f ? 0.0f : f; };  // fold -0 into +0
    uint32_t b[3];
    const float v[3] = { norm(x), norm(y), norm(z) };
    std::memcpy(b, v, sizeof(b));
    const auto key = std::make_tuple(b[0], b[1], b[2]);
    auto it = index_.find(key);
    if (it != index_.end())
      return it->second;
    const int idx = static_cast<int>(verts.size());
    verts.push_back(Vector3d{ static_cast<double>(v[0]), static_cast<double>(v[1]), static_cast<double>(v[2]) });
    index_.emplace(key, idx);
    return idx;
  }

private:
  std::map<std::tuple<uint32_t, uint32_t, uint32_t>, int> index_;
};

inline Mesh::Ptr loadBinaryStl(const std::vector<uint8_t>& d)
{
  if (d.size() < 84)
    return nullptr;
  uint32_t n = 0;
  std::memcpy(&n, d.data() + 80, 4);
  if (d.size() < 84 + static_cast<std::size_t>(n) * 50)
    return nullptr;
  auto verts = std::make_shared<VectorVector3d>();
  auto tris = std::make_shared<VectorXi>();
  tris->reserve(static_cast<std::size_t>(n) * 4);
  VertexWelder welder;
  for (uint32_t t = 0; t < n; t++)
  {
    const uint8_t* rec = d.data() + 84 + static_cast<std::size_t>(t) * 50 + 12;  // skip the facet normal
    tris->push_back(3);
    for (int j = 0; j < 3; j++)
    {
      float v[3];
      std::memcpy(v, rec + 12 * j, 12);
      tris->push_back(welder.add(v[0], v[1], v[2], *verts));
    }
  }
  return std::make_shared<Mesh>(verts, tris);
}

inline Mesh::Ptr loadAsciiPly(const std::vector<uint8_t>& d)
{
  std::istringstream in(std::string(d.begin(), d.end()));
  std::string line;
  std::size_t n_vertex = 0, n_face = 0;
  bool in_vertex = false, ascii = false;
  int vprops = 0;
  if (!std::getline(in, line) || line.compare(0, 3, "ply") != 0)
    return nullptr;
  while (std::getline(in, line))
  {
    if (!line.empty() && line.back() == '\r')
      line.pop_back();
    std::istringstream ls(line);
    std::string a, b;
    ls >> a >> b;
    if (a == "format" && b == "ascii")
      ascii = true;
    else if (a == "element")
    {
      in_vertex = (b == "vertex");
      std::size_t cnt = 0;
      ls >> cnt;
      if (b == "vertex")
        n_vertex = cnt;
      else if (b == "face")
        n_face = cnt;
    }
    else if (a == "property" && in_vertex)
      vprops++;
    else if (a == "end_header")
      break;
  }
  if (!ascii || vprops < 3)
    return nullptr;
  auto verts = std::make_shared<VectorVector3d>();
  auto tris = std::make_shared<VectorXi>();
  VertexWelder welder;
  std::vector<int> remap(n_vertex);
  for (std::size_t i = 0; i < n_vertex; i++)
  {
    if (!std::getline(in, line))
      return nullptr;
    std::istringstream ls(line);
    float x, y, z;
    if (!(ls >> x >> y >> z))
      return nullptr;
    remap[i] = welder.add(x, y, z, *verts);
  }
  for (std::size_t i = 0; i < n_face; i++)
  {
    if (!std::getline(in, line))
      return nullptr;
    std::istringstream ls(line);
    int cnt = 0;
    ls >> cnt;
    std::vector<int> idx(static_cast<std::size_t>(cnt > 0 ? cnt : 0));
    for (int j = 0; j < cnt; j++)
      ls >> idx[static_cast<std::size_t>(j)];
    for (int j = 1; j + 1 < cnt; j++)  // fan triangulation
    {
      tris->push_back(3);
      tris->push_back(remap[static_cast<std::size_t>(idx[0])]);
      tris->push_back(remap[static_cast<std::size_t>(idx[static_cast<std::size_t>(j)])]);
      tris->push_back(remap[static_cast<std::size_t>(idx[static_cast<std::size_t>(j) + 1])]);
    }
  }
  return std::make_shared<Mesh>(verts, tris);
}

/** @brief First mesh of a file: binary STL or ASCII PLY, triangulated, identical vertices welded. */
inline Mesh::Ptr createMeshFromPath(const std::string& path)
{
  std::vector<uint8_t> d;
  if (!readFile(path, d))
    return nullptr;
  if (d.size() >= 3 && std::memcmp(d.data(), "ply", 3) == 0)
    return loadAsciiPly(d);
  return loadBinaryStl(d);
}

/** @brief tesseract_collision::writeSimplePlyFile: ASCII PLY with optional per-vertex uchar colours. */
inline bool writeSimplePlyFile(const std::string& path, const VectorVector3d& vertices, const std::vector<Vector3i>& colors,
                               const VectorXi& faces, int num_faces)
{
  std::ofstream f(path);
  if (f.fail())
    return false;
  f << "ply\nformat ascii 1.0\ncomment made by tesseract\n";
  f << "element vertex " << vertices.size() << "\nproperty float x\nproperty float y\nproperty float z\n";
  if (!colors.empty())
    f << "property uchar red\nproperty uchar green\nproperty uchar blue\n";
  f << "element face " << num_faces << "\nproperty list uchar int vertex_indices\nend_header\n";
  f << std::fixed << std::setprecision(std::numeric_limits<float>::digits10 + 1);
  for (std::size_t i = 0; i < vertices.size(); i++)
  {
    f << vertices[i][0] << " " << vertices[i][1] << " " << vertices[i][2];
    if (!colors.empty())
    {
      const Vector3i& c = colors.size() == 1 ? colors[0] : colors[i];
      f << " " << c[0] << " " << c[1] << " " << c[2];
    }
    f << "\n";
  }
  std::size_t idx = 0;
  for (int i = 0; i < num_faces; i++)
  {
    const int nvert = faces[idx];
    for (int j = 0; j < nvert; j++)
      f << faces[idx++] << " ";
    f << faces[idx++] << "\n";
  }
  return !f.fail();
}

// ---- octomap .bt (binary tree) reader: occupied leaves as (min-corner key, depth) -----------------------------------------
struct BtLeaf
{
  uint16_t kx, ky, kz;
  uint8_t depth;  // 16 = finest
};

struct BtTree
{
  double resolution = 0.0;
  std::vector<BtLeaf> occupied;
};

namespace detail
{
inline bool readBtNode(const std::vector<uint8_t>& d, std::size_t& pos, uint32_t x, uint32_t y, uint32_t z, int depth, BtTree& out)
{
  if (pos + 2 > d.size() || depth >= 16)
    return false;
  const unsigned bits = d[pos] | (static_cast<unsigned>(d[pos + 1]) << 8);
  pos += 2;
  const uint32_t half = 1u << (15 - depth);
  for (int i = 0; i < 8; i++)
  {
    const unsigned c = (bits >> (2 * i)) & 3u;  // 2 bits per child: 1 free leaf, 2 occupied leaf, 3 inner node, 0 unknown
    if (c == 2u)
      out.occupied.push_back(BtLeaf{ static_cast<uint16_t>(x + ((i & 1) ? half : 0)), static_cast<uint16_t>(y + ((i & 2) ? half : 0)),
                                     static_cast<uint16_t>(z + ((i & 4) ? half : 0)), static_cast<uint8_t>(depth + 1) });
  }
  for (int i = 0; i < 8; i++)
    if (((bits >> (2 * i)) & 3u) == 3u)
      if (!readBtNode(d, pos, x + ((i & 1) ? half : 0), y + ((i & 2) ? half : 0), z + ((i & 4) ? half : 0), depth + 1, out))
        return false;
  return true;
}
}  // namespace detail

inline bool readBt(const std::string& path, BtTree& out)
{
  std::vector<uint8_t> d;
  if (!readFile(path, d))
    return false;
  std::size_t pos = 0;
  bool have_data = false;
  while (pos < d.size() && !have_data)
  {
    std::size_t e = pos;
    while (e < d.size() && d[e] != '\n')
      e++;
    const std::string line(d.begin() + static_cast<std::ptrdiff_t>(pos), d.begin() + static_cast<std::ptrdiff_t>(e));
    pos = e + 1;
    if (line.compare(0, 4, "res ") == 0)
      out.resolution = std::atof(line.c_str() + 4);
    else if (line.compare(0, 4, "data") == 0)
      have_data = true;
  }
  if (!have_data || out.resolution <= 0.0)
    return false;
  if (pos >= d.size())
    return true;  // empty tree
  return detail::readBtNode(d, pos, 0, 0, 0, 0, out);
}
}  // namespace pcs_compat
