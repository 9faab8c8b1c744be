// detection_unit.cpp -- the reference's pcs_detection gtests (pcs_detection/test/utils_unit.cpp:38-64,
// pcs_detection/test/point_cloud_annotator_unit.cpp:61-90) re-expressed against the drop-in adapters, plus the
// detector and the fused annotator mode.  Runs on a CUDA device (everything goes through libpcs_b200's C ABI).
#include <pcs_detection/hsv_thresholding.h>
#include <pcs_detection/point_cloud_annotator.h>
#include <pcs_detection/utils.h>
#include <pcs_ros/point_cloud_annotator_node.h>

#include "mini_test.h"

using namespace pcs_detection;

static PointCloud<PointXYZRGB>::Ptr rampCloud()
{
  // 255 x 1 cloud, x = i / 100, grey ramp (utils_unit.cpp:16-35)
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

TEST(UtilsUnit, CloudImageManipulation)
{
  auto cloud = rampCloud();
  auto position_image = std::make_shared<Mat>();
  auto image_2d = std::make_shared<Mat>();
  cloudToImage(cloud, *position_image, *image_2d);
  EXPECT_EQ(position_image->rows, image_2d->rows);
  EXPECT_EQ(position_image->cols, image_2d->cols);
  PointCloud<PointXYZRGB>::Ptr output_cloud = imageToCloud(*image_2d, *position_image, cloud->header);
  EXPECT_EQ(cloud->points.size(), output_cloud->points.size());
  EXPECT_EQ(cloud->width, output_cloud->width);
  EXPECT_EQ(cloud->height, output_cloud->height);
  EXPECT_EQ(cloud->is_dense, output_cloud->is_dense);
  for (std::size_t i = 0; i < cloud->points.size(); i++)
  {
    EXPECT_EQ(cloud->points[i].x, output_cloud->points[i].x);
    EXPECT_EQ(cloud->points[i].y, output_cloud->points[i].y);
    EXPECT_EQ(cloud->points[i].z, output_cloud->points[i].z);
    EXPECT_EQ(cloud->points[i].r, output_cloud->points[i].r);
    EXPECT_EQ(cloud->points[i].g, output_cloud->points[i].g);
    EXPECT_EQ(cloud->points[i].b, output_cloud->points[i].b);
  }
}

TEST(HsvThresholdingUnit, detector)
{
  // a 4 x 64 image: left half the painted colour (B,G,R) = (140,90,50) -> H 107, S 164, V 140: inside the thresholds
  Mat img(4, 64, pcs_compat::CV_8UC3);
  for (int y = 0; y < 4; y++)
    for (int x = 0; x < 64; x++)
    {
      uint8_t* px = img.data + img.step * y + 3 * x;
      px[0] = x < 32 ? 140 : 10;
      px[1] = x < 32 ? 90 : 200;
      px[2] = x < 32 ? 50 : 30;
    }
  Mat mask, inv;
  EXPECT_TRUE(hsvThresholdingDetector(img, mask));
  EXPECT_TRUE(hsvThresholdingDetector(img, inv, true));
  EXPECT_EQ(mask.rows, 4);
  EXPECT_EQ(mask.cols, 64);
  for (int y = 0; y < 4; y++)
    for (int x = 0; x < 64; x++)
    {
      EXPECT_EQ(mask.at<uint8_t>(y, x), x < 32 ? 255 : 0);
      EXPECT_EQ(inv.at<uint8_t>(y, x), x < 32 ? 0 : 255);
    }
}

static std::vector<Mat> ImageAnnotatorCallback_ReturnInputs(const std::vector<Mat> input_images) { return input_images; }
static std::vector<Mat> ImageAnnotatorCallback_Throw(const std::vector<Mat>) { throw std::exception(); }
static int g_results = 0;
static void ResultsCallback_ReturnCleanly(pointCloudVec results) { g_results += static_cast<int>(results.size()); }
static void ResultsCallback_Throw(pointCloudVec) { throw std::exception(); }

TEST(PointCloudAnnotatorUnit, Construction)
{
  PointCloudAnnotator pca1(&ImageAnnotatorCallback_ReturnInputs, &ResultsCallback_ReturnCleanly);
  PointCloudAnnotator pca2(&ImageAnnotatorCallback_Throw, &ResultsCallback_Throw, 123);
}

TEST(PointCloudAnnotatorUnit, addPointCloud)
{
  auto cloud = rampCloud();
  {
    g_results = 0;
    PointCloudAnnotator pca1(&ImageAnnotatorCallback_ReturnInputs, &ResultsCallback_ReturnCleanly, 2);
    EXPECT_TRUE(pca1.addPointCloud(cloud));
    EXPECT_EQ(g_results, 0);
    EXPECT_TRUE(pca1.addPointCloud(cloud));
    EXPECT_EQ(g_results, 2);
  }
  {
    // fails as soon as the buffer gets big enough to trigger annotateImages
    PointCloudAnnotator pca1(&ImageAnnotatorCallback_Throw, &ResultsCallback_ReturnCleanly, 3);
    EXPECT_TRUE(pca1.addPointCloud(cloud));
    EXPECT_TRUE(pca1.addPointCloud(cloud));
    EXPECT_FALSE(pca1.addPointCloud(cloud));
  }
  {
    PointCloudAnnotator pca1(&ImageAnnotatorCallback_ReturnInputs, &ResultsCallback_Throw, 3);
    EXPECT_TRUE(pca1.addPointCloud(cloud));
    EXPECT_TRUE(pca1.addPointCloud(cloud));
    EXPECT_FALSE(pca1.addPointCloud(cloud));
  }
}

static pointCloudVec g_last;
static void ResultsCallback_Keep(pointCloudVec results) { g_last = results; }
static std::vector<Mat> ImageAnnotatorCallback_Hsv(const std::vector<Mat> input_images)
{
  // what pcs_ros/src/hsv_thresholding_node.cpp:8-28 answers: the mask replicated to three channels
  std::vector<Mat> out;
  for (const Mat& img : input_images)
  {
    Mat m3(img.rows, img.cols, pcs_compat::CV_8UC3);
    pcs_hsv_params p;
    pcs_hsv_default_params(&p);
    if (pcs_hsv_annotation_image(pcs_compat::default_ctx(), img.data, img.cols, img.rows, img.step, &p, m3.data) != PCS_OK)
      throw std::runtime_error(pcs_last_error());
    out.push_back(m3);
  }
  return out;
}

TEST(PointCloudAnnotatorUnit, fusedModeEqualsCallbackMode)
{
  PointCloud<PointXYZRGB>::Ptr cloud(new PointCloud<PointXYZRGB>());
  cloud->width = 64;
  cloud->height = 48;
  cloud->is_dense = false;
  cloud->points.resize(64 * 48);
  for (size_t i = 0; i < cloud->points.size(); ++i)
  {
    PointXYZRGB& p = cloud->points[i];
    p.x = static_cast<float>(i % 64) * 0.01f;
    p.y = static_cast<float>(i / 64) * 0.01f;
    p.z = (i % 17 == 0) ? std::numeric_limits<float>::quiet_NaN() : 0.5f;
    const bool painted = (i / 7) % 3 == 0;
    p.b = painted ? 140 : static_cast<uint8_t>(i * 31);
    p.g = painted ? 90 : static_cast<uint8_t>(i * 17);
    p.r = painted ? 50 : static_cast<uint8_t>(i * 5);
  }
  PointCloudAnnotator slow(&ImageAnnotatorCallback_Hsv, &ResultsCallback_Keep, 2);
  EXPECT_TRUE(slow.addPointCloud(cloud));
  EXPECT_TRUE(slow.addPointCloud(cloud));
  pointCloudVec a = g_last;
  PointCloudAnnotator fast(&ImageAnnotatorCallback_Throw, &ResultsCallback_Keep, 2);  // the callback is never used
  pcs_annotate_params prm;
  pcs_annotate_default_params(&prm);
  fast.useFusedHsvDetector(prm);
  EXPECT_TRUE(fast.addPointCloud(cloud));
  EXPECT_TRUE(fast.addPointCloud(cloud));
  pointCloudVec b = g_last;
  EXPECT_EQ(a.size(), 2u);
  EXPECT_EQ(b.size(), 2u);
  for (size_t f = 0; f < 2 && f < a.size() && f < b.size(); f++)
  {
    EXPECT_EQ(a[f]->points.size(), b[f]->points.size());
    EXPECT_EQ(a[f]->width, b[f]->width);
    EXPECT_TRUE(std::memcmp(a[f]->points.data(), b[f]->points.data(), a[f]->points.size() * sizeof(PointXYZRGB)) == 0);
    int detected = 0;
    for (const PointXYZRGB& p : b[f]->points)
      detected += p.r == 255;
    EXPECT_TRUE(detected > 500 && detected < 2500);
  }
}

TEST(PointCloudAnnotatorUnit, fusedModeRefusesUnorganizedClouds)
{
  // the fused path skips cloudToImage, so its organized-cloud check (utils.h:103-105) is made at addPointCloud
  PointCloudAnnotator fast(&ImageAnnotatorCallback_Throw, &ResultsCallback_Keep, 1);
  pcs_annotate_params prm;
  pcs_annotate_default_params(&prm);
  fast.useFusedHsvDetector(prm);
  PointCloud<PointXYZRGB>::Ptr bad(new PointCloud<PointXYZRGB>());
  bad->width = 64;
  bad->height = 48;
  bad->points.resize(100);  // width * height != points.size()
  EXPECT_FALSE(fast.addPointCloud(bad));
  PointCloud<PointXYZRGB>::Ptr one(new PointCloud<PointXYZRGB>());
  one->width = 1;
  one->height = 1;
  one->points.resize(1);
  EXPECT_FALSE(fast.addPointCloud(one));
  g_last.clear();
  PointCloud<PointXYZRGB>::Ptr ok(new PointCloud<PointXYZRGB>());
  ok->width = 8;
  ok->height = 4;
  ok->points.resize(32);
  EXPECT_TRUE(fast.addPointCloud(ok));  // the refused clouds were never queued
  EXPECT_EQ(g_last.size(), 1u);
}

TEST(UtilsUnit, applyMask)
{
  // utils.h:82-86: input.mul(mask), saturated
  Mat img(2, 3, pcs_compat::CV_8UC3), mask(2, 3, pcs_compat::CV_8UC3), out;
  for (int i = 0; i < 18; i++)
  {
    img.data[i] = static_cast<uint8_t>(10 * i);
    mask.data[i] = static_cast<uint8_t>(i % 3);  // 0, 1, 2
  }
  EXPECT_TRUE(pcs_detection::applyMask(img, mask, out));
  EXPECT_EQ(out.rows, 2);
  EXPECT_EQ(out.cols, 3);
  for (int i = 0; i < 18; i++)
  {
    const unsigned v = 10u * i * (i % 3);
    EXPECT_EQ(static_cast<unsigned>(out.data[i]), v > 255u ? 255u : v);
  }
}

TEST(PointCloudAnnotatorNodeUnit, toggleAndPublish)
{
  // pcs_ros/src/point_cloud_annotator_node.cpp:57-61,106-115,157-163 without ROS
  auto cloud = rampCloud();
  cloud->width = 51;
  cloud->height = 5;  // 255 points as a 51 x 5 organized cloud
  for (size_t i = 0; i < cloud->points.size(); ++i)
    if (i % 3 == 0)
    {
      cloud->points[i].b = 140;
      cloud->points[i].g = 90;
      cloud->points[i].r = 50;
    }
  int published = 0, debug_images = 0, detected = 0;
  pcs_ros::PointCloudAnnotatorNode node([&](PointCloud<PointXYZRGB>::ConstPtr c) {
    published++;
    detected = 0;
    for (const PointXYZRGB& p : c->points)
      detected += p.r == 255 && p.g == 255 && p.b == 255;
  });
  node.setDebugImageSink([&](const Mat&) { debug_images++; });
  node.subscriberCallback(cloud);
  EXPECT_EQ(published, 1);
  EXPECT_EQ(debug_images, 1);
  EXPECT_EQ(detected, 85);
  EXPECT_TRUE(node.toggleAnnotation(false) == "Annotation disabled");
  node.subscriberCallback(cloud);
  EXPECT_EQ(published, 1);  // dropped while disabled
  EXPECT_TRUE(node.toggleAnnotation(true) == "Annotation enabled");
  pcs_annotate_params prm;
  pcs_annotate_default_params(&prm);
  node.setFusedHsvDetector(prm);
  node.subscriberCallback(cloud);
  EXPECT_EQ(published, 2);
  EXPECT_EQ(debug_images, 1);  // the fused path never materialises the 2D annotation
  EXPECT_EQ(detected, 85);
}

int main() { return mini_test::run_all(); }
