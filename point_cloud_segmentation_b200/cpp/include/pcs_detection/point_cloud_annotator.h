// pcs_detection/point_cloud_annotator.h -- drop-in for pcs_detection/include/pcs_detection/point_cloud_annotator.h:39-103
// and pcs_detection/src/pcs_detection/point_cloud_annotator.cpp:7-82.
//
// Same class, constructor and methods.  Two execution modes:
//   * callback mode (the reference's): addPointCloud -> cloudToImage, queue; at batch_size call the user's detector on
//     the batch of 2D images, imageToCloud per item, deliver through results_callback.
//   * fused mode (useFusedHsvDetector): when the detector is the HSV threshold, the whole batch goes through one
//     pcs_annotate_frames call -- no position / BGR image is built, the service hop of
//     pcs_ros/src/point_cloud_annotator_node.cpp:68-99 disappears, and an optional voxel map is updated in the same pass.
// Deliberate deviation: the reference fills the detector batch with the FRONT frame batch_size times
// (point_cloud_annotator.cpp:38-42); here the batch holds the distinct queued frames.  Identical at the node's
// batch_size = 1.  Exception behaviour is the reference's: a throwing detector pops one entry and returns false
// (:47-58), a throwing results callback returns false after the batch was consumed (:72-80).
#ifndef PCS_DETECTION_POINT_CLOUD_ANNOTATOR_H
#define PCS_DETECTION_POINT_CLOUD_ANNOTATOR_H

#include <cassert>
#include <deque>
#include <functional>
#include <mutex>

#include "pcs_compat/types.h"
#include "pcs_detection/utils.h"

namespace pcs_detection
{
typedef std::vector<PointCloud<PointXYZRGB>::ConstPtr> pointCloudVec;

/** @brief Everything needed to process a cloud later (reference :46-62). */
struct PointCloudData
{
  PointCloudData() = default;
  PointCloudData(PointCloud<PointXYZRGB>::ConstPtr& cloud, std::shared_ptr<Mat>& position_image, std::shared_ptr<Mat>& image_2d)
    : cloud_(cloud), position_image_(position_image), image_2d_(image_2d)
  {
  }
  PointCloud<PointXYZRGB>::ConstPtr cloud_;
  std::shared_ptr<Mat> position_image_;
  std::shared_ptr<Mat> image_2d_;
};

class PointCloudAnnotator
{
public:
  PointCloudAnnotator(std::function<std::vector<Mat>(const std::vector<Mat>)> image_annotator_callback,
                      std::function<void(pointCloudVec)> results_callback, long unsigned int batch_size = 1)
    : image_annotator_callback_(std::move(image_annotator_callback))
    , results_callback_(std::move(results_callback))
    , batch_size_(batch_size)
  {
    assert(batch_size_ <= 1024);  // sanity check of the reference (:78)
  }

  /** @brief Route batches through the fused GPU path with the HSV detector; `map` (may be NULL) receives the labelled
   * points of every frame with the given colorPassthrough filter. */
  void useFusedHsvDetector(const pcs_annotate_params& params, pcs_map* map = nullptr)
  {
    fused_ = true;
    fused_params_ = params;
    fused_map_ = map;
  }

  /** @brief Adds a cloud to the queue; annotates when the batch is full.  @return false if annotation failed. */
  bool addPointCloud(PointCloud<PointXYZRGB>::ConstPtr input_cloud)
  {
    auto position_image = std::make_shared<Mat>();
    auto image_2d = std::make_shared<Mat>();
    if (!fused_)
      cloudToImage(input_cloud, *position_image, *image_2d);
    else if (!input_cloud || (input_cloud->width == 1 && input_cloud->height == 1) ||
             input_cloud->points.size() != static_cast<std::size_t>(input_cloud->width) * input_cloud->height)
    {
      // the fused path skips cloudToImage and with it the reference's organized-cloud check (utils.h:103-105); the
      // kernel reads width * height points per frame, so anything else is refused here instead of asserted there
      pcs_compat::logError("addPointCloud: cloud is not organized (width * height != points.size(), or 1 x 1)");
      return false;
    }
    PointCloudData data(input_cloud, position_image, image_2d);
    {
      std::lock_guard<std::mutex> lock(buffer_mutex_);
      input_buffer_.push_back(data);
    }
    if (bufferSize() >= batch_size_)
    {
      if (!annotateImages())
      {
        pcs_compat::logError("Annotate Images failed");
        return false;
      }
    }
    return true;
  }

  /** @brief Takes one batch from the queue, annotates it and delivers the results.  @return false if failed. */
  bool annotateImages()
  {
    if (bufferSize() < batch_size_)
      return false;
    return fused_ ? annotateFused() : annotateWithCallback();
  }

protected:
  std::size_t bufferSize()
  {
    std::lock_guard<std::mutex> lock(buffer_mutex_);  // the reference reads size() unlocked (Q3)
    return input_buffer_.size();
  }

  bool annotateWithCallback()
  {
    std::vector<Mat> vec(batch_size_);
    {
      std::lock_guard<std::mutex> lock(buffer_mutex_);
      for (std::size_t idx = 0; idx < batch_size_; idx++)
        vec[idx] = *input_buffer_[idx].image_2d_;
    }
    std::vector<Mat> image_annotations;
    try
    {
      image_annotations = image_annotator_callback_(vec);
    }
    catch (...)
    {
      pcs_compat::logError("Image Annotator Callback Exception");
      std::lock_guard<std::mutex> lock(buffer_mutex_);
      input_buffer_.pop_front();
      return false;
    }
    pointCloudVec results(batch_size_);
    try
    {
      std::lock_guard<std::mutex> lock(buffer_mutex_);
      for (std::size_t idx = 0; idx < batch_size_; idx++)
      {
        results[idx] = imageToCloud(image_annotations.at(idx), *input_buffer_.front().position_image_,
                                    input_buffer_.front().cloud_->header);
        input_buffer_.pop_front();
      }
    }
    catch (...)
    {
      pcs_compat::logError("Applying annotations failed");
      return false;
    }
    return deliver(results);
  }

  bool annotateFused()
  {
    // the batch stays queued until every call has succeeded; on failure only the front entry is dropped, which is what
    // the reference does when the detector callback fails (point_cloud_annotator.cpp:51-56)
    std::vector<PointCloudData> batch;
    {
      std::lock_guard<std::mutex> lock(buffer_mutex_);
      for (std::size_t idx = 0; idx < batch_size_; idx++)
        batch.push_back(input_buffer_[idx]);
    }
    auto drop = [this](std::size_t n) {
      std::lock_guard<std::mutex> lock(buffer_mutex_);
      for (std::size_t idx = 0; idx < n && !input_buffer_.empty(); idx++)
        input_buffer_.pop_front();
    };
    pointCloudVec results(batch_size_);
    // frames of one shape go through one call; a batch of mixed shapes falls back to one call per frame
    std::size_t i = 0;
    while (i < batch.size())
    {
      std::size_t j = i + 1;
      while (j < batch.size() && batch[j].cloud_->width == batch[i].cloud_->width && batch[j].cloud_->height == batch[i].cloud_->height)
        j++;
      const std::size_t nf = j - i, ppf = batch[i].cloud_->points.size();
      std::vector<PointXYZRGB> in(nf * ppf), out(nf * ppf);
      for (std::size_t f = 0; f < nf; f++)
        std::memcpy(&in[f * ppf], batch[i + f].cloud_->points.data(), ppf * sizeof(PointXYZRGB));
      const int rc = pcs_annotate_frames(pcs_compat::default_ctx(), fused_map_, reinterpret_cast<const pcs_point_xyzrgb*>(in.data()),
                                         static_cast<int>(nf), static_cast<int>(batch[i].cloud_->width),
                                         static_cast<int>(batch[i].cloud_->height), nullptr, &fused_params_,
                                         reinterpret_cast<pcs_point_xyzrgb*>(out.data()), nullptr);
      if (rc != PCS_OK)
      {
        pcs_compat::logError(pcs_last_error());
        drop(1);
        return false;
      }
      for (std::size_t f = 0; f < nf; f++)
      {
        PointCloud<PointXYZRGB>::Ptr c(new PointCloud<PointXYZRGB>());
        c->header = batch[i + f].cloud_->header;
        c->width = batch[i + f].cloud_->width;
        c->height = batch[i + f].cloud_->height;
        c->is_dense = false;
        c->points.assign(out.begin() + static_cast<std::ptrdiff_t>(f * ppf), out.begin() + static_cast<std::ptrdiff_t>((f + 1) * ppf));
        results[i + f] = c;
      }
      i = j;
    }
    drop(batch_size_);
    return deliver(results);
  }

  bool deliver(const pointCloudVec& results)
  {
    try
    {
      results_callback_(results);
    }
    catch (...)
    {
      pcs_compat::logError("Results Callback Exception");
      return false;
    }
    return true;
  }

  std::function<std::vector<Mat>(const std::vector<Mat>)> image_annotator_callback_;
  std::function<void(pointCloudVec)> results_callback_;
  long unsigned int batch_size_;
  std::deque<PointCloudData> input_buffer_;
  bool fused_ = false;
  pcs_annotate_params fused_params_{};
  pcs_map* fused_map_ = nullptr;

private:
  std::mutex buffer_mutex_;
};
}  // namespace pcs_detection
#endif
