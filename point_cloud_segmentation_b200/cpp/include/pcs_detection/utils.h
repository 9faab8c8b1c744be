// pcs_detection/utils.h -- drop-in for pcs_detection/include/pcs_detection/utils.h:42-171.
// cloudToImage / imageToCloud keep their signatures; the per-pixel loops run on the GPU (pcs_cloud_to_image,
// pcs_image_to_cloud).  The reference's asserts (:103-105, :142-143) become exceptions so that release builds still
// notice them.
#ifndef PCS_DETECTION_UTILS_H
#define PCS_DETECTION_UTILS_H

#include <limits>
#include <stdexcept>
#include <string>

#include "pcs_compat/types.h"

namespace pcs_detection
{
using pcs_compat::Mat;
using pcs_compat::PCLHeader;
using pcs_compat::PointCloud;
using pcs_compat::PointXYZRGB;

/** @brief Human readable name of a matrix type, e.g. "8UC3" (reference :42-81). */
inline std::string type2str(const int type)
{
  switch (type)
  {
    case pcs_compat::CV_8UC1:
      return "8UC1";
    case pcs_compat::CV_8UC3:
      return "8UC3";
    case pcs_compat::CV_64FC3:
      return "64FC3";
    default:
      return "UserC1";
  }
}

/** @brief masked_image = input_image.mul(mask): per-element product, saturated to the element type (reference :82-86).
 * Host-side helper exactly as in the reference (nothing on the hot path calls it); both images must have the same size
 * and type, as cv::Mat::mul requires (it throws otherwise, so does this). */
inline bool applyMask(const Mat& input_image, const Mat& mask, Mat& masked_image)
{
  if (input_image.rows != mask.rows || input_image.cols != mask.cols || input_image.type() != mask.type())
    throw std::invalid_argument("applyMask: size or type mismatch (cv::Mat::mul)");
  Mat out(input_image.rows, input_image.cols, input_image.type());
  const std::size_t row_elems = static_cast<std::size_t>(input_image.cols) * (input_image.type() == pcs_compat::CV_8UC1 ? 1 : 3);
  for (int r = 0; r < input_image.rows; r++)
  {
    if (input_image.type() == pcs_compat::CV_64FC3)
    {
      const double* a = reinterpret_cast<const double*>(input_image.data + input_image.step * static_cast<std::size_t>(r));
      const double* b = reinterpret_cast<const double*>(mask.data + mask.step * static_cast<std::size_t>(r));
      double* o = reinterpret_cast<double*>(out.data + out.step * static_cast<std::size_t>(r));
      for (std::size_t i = 0; i < row_elems; i++)
        o[i] = a[i] * b[i];
    }
    else
    {
      const uint8_t* a = input_image.data + input_image.step * static_cast<std::size_t>(r);
      const uint8_t* b = mask.data + mask.step * static_cast<std::size_t>(r);
      uint8_t* o = out.data + out.step * static_cast<std::size_t>(r);
      for (std::size_t i = 0; i < row_elems; i++)
      {
        const unsigned v = static_cast<unsigned>(a[i]) * b[i];
        o[i] = static_cast<uint8_t>(v > 255u ? 255u : v);  // cv::saturate_cast<uchar>
      }
    }
  }
  masked_image = out;
  return true;
}

/** @brief Organized XYZRGB cloud -> 64FC3 position image + 8UC3 BGR image (reference :98-129). */
inline void cloudToImage(const PointCloud<PointXYZRGB>::ConstPtr cloud, Mat& position_image, Mat& color_image)
{
  if (cloud->width == 1 && cloud->height == 1)
    throw std::invalid_argument("cloudToImage: cloud is not organized");  // reference assert :103
  if (static_cast<std::size_t>(cloud->width) * cloud->height != cloud->points.size())
    throw std::invalid_argument("cloudToImage: width * height != points.size()");
  position_image = Mat(static_cast<int>(cloud->height), static_cast<int>(cloud->width), pcs_compat::CV_64FC3);
  color_image = Mat(static_cast<int>(cloud->height), static_cast<int>(cloud->width), pcs_compat::CV_8UC3);
  const int rc = pcs_cloud_to_image(pcs_compat::default_ctx(), reinterpret_cast<const pcs_point_xyzrgb*>(cloud->points.data()),
                                    static_cast<int>(cloud->width), static_cast<int>(cloud->height),
                                    reinterpret_cast<double*>(position_image.data), color_image.data);
  if (rc != PCS_OK)
    throw std::runtime_error(pcs_last_error());
}

/** @brief 3-channel annotation + position image -> XYZRGB cloud coloured by the annotation (reference :138-171). */
inline PointCloud<PointXYZRGB>::Ptr imageToCloud(const Mat& color_image, const Mat& position_image,
                                                 const PCLHeader& header = PCLHeader())
{
  if (color_image.rows != position_image.rows || color_image.cols != position_image.cols)
    throw std::invalid_argument("imageToCloud: image sizes differ");  // reference asserts :142-143
  PointCloud<PointXYZRGB>::Ptr cloud(new PointCloud<PointXYZRGB>());
  cloud->header = header;
  cloud->points.resize(static_cast<std::size_t>(color_image.rows) * color_image.cols);
  const int rc = pcs_image_to_cloud(pcs_compat::default_ctx(), color_image.data, reinterpret_cast<const double*>(position_image.data),
                                    color_image.cols, color_image.rows, reinterpret_cast<pcs_point_xyzrgb*>(cloud->points.data()));
  if (rc != PCS_OK)
    throw std::runtime_error(pcs_last_error());
  cloud->height = static_cast<uint32_t>(color_image.rows);
  cloud->width = static_cast<uint32_t>(color_image.cols);
  cloud->is_dense = false;  // reference :169
  return cloud;
}
}  // namespace pcs_detection
#endif
