// octomap_mesh_masking_unit.cpp -- the reference's pcs_scan_integration gtests
// (pcs_scan_integration/test/octomap_mesh_masking_unit.cpp:45-185) re-expressed against the drop-in adapter.
// DATA_DIR comes from $PCS_TEST_DATA_DIR and must hold plane_4m.stl, box_2m.ply and box_2m.bt (the pytest wrapper
// regenerates them from tests/golden; the .bt is byte-identical to the reference fixture).
#include <pcs_scan_integration/octomap_mesh_masking.h>

#include "mini_test.h"

using namespace pcs_scan_integration;

static std::string DATA_DIR()
{
  const char* d = std::getenv("PCS_TEST_DATA_DIR");
  return d ? d : ".";
}

static PointCloud<PointXYZRGB>::Ptr rampCloud()
{
  PointCloud<PointXYZRGB>::Ptr cloud(new PointCloud<PointXYZRGB>());
  cloud->width = 255;
  cloud->height = 1;
  cloud->is_dense = false;
  cloud->points.resize(cloud->width * cloud->height);
  for (size_t i = 0; i < cloud->points.size(); ++i)
  {
    cloud->points[i].x = static_cast<float>(i) / 100.f;
    cloud->points[i].y = 0.0;
    cloud->points[i].z = 0.0;
    cloud->points[i].r = static_cast<uint8_t>(i);
    cloud->points[i].g = static_cast<uint8_t>(i);
    cloud->points[i].b = static_cast<uint8_t>(i);
  }
  return cloud;
}

TEST(OctomapMeshMaskUnit, getSet)
{
  auto cloud = rampCloud();
  OctomapMeshMask masker;
  EXPECT_TRUE(masker.getOctree() == nullptr);
  EXPECT_TRUE(masker.getInputMesh() == nullptr);
  EXPECT_TRUE(masker.getMaskedMesh() == nullptr);
  {
    auto octree = std::make_shared<Octree>(*cloud, 0.1, Octree::SubType::BOX, true);
    masker.setOctree(octree);
    EXPECT_TRUE(masker.getOctree() != nullptr);
    masker.setOctree(nullptr);
    masker.setOctree(cloud, 0.1);
    EXPECT_TRUE(masker.getOctree() != nullptr);
  }
  {
    masker.setOctree(nullptr);
    masker.setOctree(cloud, 0.1, 0, 255, false);
    Octree::ConstPtr octree = masker.getOctree();
    EXPECT_TRUE(octree != nullptr);
    EXPECT_EQ(octree->occupiedVoxels(), 26u);  // x = 0 .. 2.54 m at 0.1 m: keys 32768 .. 32793
  }
  {
    std::string path = DATA_DIR() + "/plane_4m.stl";
    masker.setInputMesh(path);
    auto input_mesh = masker.getInputMesh();
    EXPECT_EQ(input_mesh->getVertices()->size(), 6273u);
    EXPECT_EQ(input_mesh->getTriangles()->size(), 12288u * 4);
  }
  {
    std::string path = DATA_DIR() + "/box_2m.ply";
    Mesh::Ptr mesh = pcs_compat::createMeshFromPath(path);
    masker.setInputMesh(mesh);
    auto input_mesh = masker.getInputMesh();
    EXPECT_EQ(input_mesh->getVertices()->size(), 8u);
    EXPECT_EQ(input_mesh->getTriangles()->size(), 12u * 4);
  }
}

TEST(OctomapMeshMaskUnit, colorPassthrough)
{
  auto cloud = rampCloud();
  for (int limit_neg = 0; limit_neg < 2; limit_neg++)
    for (int lower_lim = 0; lower_lim < 255; lower_lim += 25)
      for (int upper_lim = lower_lim; upper_lim < 255; upper_lim += 25)
      {
        auto result_cloud = pcs_scan_integration::colorPassthrough(cloud, lower_lim, upper_lim, limit_neg != 0);
        if (!limit_neg)
        {
          EXPECT_EQ(result_cloud->size(), static_cast<size_t>(upper_lim - lower_lim + 1));
          for (const PointXYZRGB& point : result_cloud->points)
          {
            EXPECT_LE(point.r, upper_lim);
            EXPECT_GE(point.b, lower_lim);
          }
        }
        else
        {
          EXPECT_EQ(result_cloud->size(), cloud->size() - static_cast<size_t>(upper_lim - lower_lim) + (upper_lim == lower_lim ? 0 : 1));
          for (const PointXYZRGB& point : result_cloud->points)
            EXPECT_FALSE(((point.g > lower_lim) && (point.g < upper_lim)));
        }
      }
}

TEST(OctomapMeshMaskUnit, maskMesh)
{
  OctomapMeshMask masker;
  std::string path = DATA_DIR() + "/box_2m.bt";
  Octree::Ptr octree(new Octree(path));
  masker.setOctree(octree);
  {
    // .bt round trip: what was read is written back byte for byte
    const std::string copy = DATA_DIR() + "/box_2m_copy.bt";
    EXPECT_TRUE(octree->writeBinary(copy));
    std::vector<uint8_t> a, b;
    EXPECT_TRUE(pcs_compat::readFile(path, a));
    EXPECT_TRUE(pcs_compat::readFile(copy, b));
    EXPECT_TRUE(a == b);
    EXPECT_EQ(octree->occupiedVoxels(), 64u);
  }
  Mesh::Ptr mesh = pcs_compat::createMeshFromPath(DATA_DIR() + "/plane_4m.stl");
  masker.setInputMesh(mesh);
  {
    EXPECT_TRUE(masker.maskMesh(OctomapMeshMask::MaskType::RETURN_COLORIZED));
    std::string output_path = DATA_DIR() + "/test_output_RETURN_COLORIZED.ply";
    EXPECT_TRUE(masker.saveMaskedMesh(output_path));
    Mesh::Ptr results = masker.getMaskedMesh();
    EXPECT_EQ(results->getVertices()->size(), 6273u);
    EXPECT_EQ(results->getTriangles()->size(), 12288u * 4);
  }
  {
    EXPECT_TRUE(masker.maskMesh(OctomapMeshMask::MaskType::RETURN_INSIDE));
    Mesh::Ptr results = masker.getMaskedMesh();
    // reference golden: 5925 / 1975*4 = one entry per Bullet contact; here one entry per contacted triangle
    EXPECT_EQ(results->getVertices()->size(), 888u * 3);
    EXPECT_EQ(results->getTriangles()->size(), 888u * 4);
  }
  {
    EXPECT_TRUE(masker.maskMesh(OctomapMeshMask::MaskType::RETURN_OUTSIDE));
    std::string output_path = DATA_DIR() + "/test_output_RETURN_OUTSIDE.ply";
    EXPECT_TRUE(masker.saveMaskedMesh(output_path));
    Mesh::Ptr results = masker.getMaskedMesh();
    EXPECT_EQ(results->getVertices()->size(), 34200u);
    EXPECT_EQ(results->getTriangles()->size(), 11400u * 4);
  }
}

int main() { return mini_test::run_all(); }
