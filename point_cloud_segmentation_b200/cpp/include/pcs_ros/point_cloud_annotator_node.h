// pcs_ros/point_cloud_annotator_node.h -- ROS-free shim of pcs_ros::PointCloudAnnotatorNode
// (pcs_ros/src/point_cloud_annotator_node.cpp:51-185).
//
// The node's logic without roscpp: the subscriber / service / publisher endpoints become plain member functions and
// std::function sinks, so a host (a ROS 1 or ROS 2 wrapper, a test, a non-ROS pipeline) wires them to its transport.
//   subscriberCallback(cloud)        reference :57-61   forwards to the annotator while annotation is enabled
//   imageAnnotator(images)           reference :68-99   the reference ships every image to the /perform_detection
//                                                       service over TCP; here the detector is an in-process callable
//                                                       (default: the HSV detector on the GPU, mask replicated to 3
//                                                       channels like pcs_ros/src/hsv_thresholding_node.cpp:20-23)
//   publisherCallback(results)       reference :106-115 hands every annotated cloud to the output sink, in order
//                                                       (the reference also sleeps 1/60 s between clouds of a batch)
//   toggleAnnotation(enable)         reference :157-163 std_srvs/SetBool semantics: returns the message the service sends
//   setFusedHsvDetector()            no counterpart: batches skip the 2D round trip entirely (pcs_annotate_frames)
#ifndef PCS_ROS_POINT_CLOUD_ANNOTATOR_NODE_H
#define PCS_ROS_POINT_CLOUD_ANNOTATOR_NODE_H

#include <pcs_detection/hsv_thresholding.h>
#include <pcs_detection/point_cloud_annotator.h>

namespace pcs_ros
{
using pcs_detection::Mat;
using pcs_detection::PointCloud;
using pcs_detection::PointXYZRGB;

class PointCloudAnnotatorNode
{
public:
  typedef std::function<Mat(const Mat&)> Detector;                               // one 8UC3 image -> one 8UC3 annotation
  typedef std::function<void(PointCloud<PointXYZRGB>::ConstPtr)> CloudSink;      // the "output" topic
  typedef std::function<void(const Mat&)> ImageSink;                             // the debug image topic

  explicit PointCloudAnnotatorNode(CloudSink output, unsigned long batch_size = 1)
    : output_(std::move(output))
    , annotation_enabled_(true)
    , annotator_([this](std::vector<Mat> input) { return imageAnnotator(input); },
                 [this](pcs_detection::pointCloudVec results) { publisherCallback(results); }, batch_size)
  {
    detector_ = [](const Mat& image) {
      Mat annotation(image.rows, image.cols, pcs_compat::CV_8UC3);
      pcs_hsv_params p;
      pcs_hsv_default_params(&p);
      if (pcs_hsv_annotation_image(pcs_compat::default_ctx(), image.data, image.cols, image.rows, image.step, &p,
                                   annotation.data) != PCS_OK)
        throw std::runtime_error(pcs_last_error());
      return annotation;
    };
  }

  void setDetector(Detector d) { detector_ = std::move(d); }
  void setDebugImageSink(ImageSink s) { debug_sink_ = std::move(s); }
  void setFusedHsvDetector(const pcs_annotate_params& params, pcs_map* map = nullptr) { annotator_.useFusedHsvDetector(params, map); }

  void subscriberCallback(const PointCloud<PointXYZRGB>::Ptr& input_cloud)
  {
    if (annotation_enabled_)
      annotator_.addPointCloud(input_cloud);
  }

  std::vector<Mat> imageAnnotator(const std::vector<Mat> input_images)
  {
    std::vector<Mat> results(input_images.size());
    for (std::size_t idx = 0; idx < input_images.size(); idx++)
    {
      results[idx] = detector_(input_images[idx]);
      if (debug_sink_)
        debug_sink_(results[idx]);
    }
    return results;
  }

  void publisherCallback(pcs_detection::pointCloudVec results)
  {
    for (auto& result : results)
      output_(result);
  }

  /** @brief SetBool service body: enables / disables annotation, returns the response message. */
  std::string toggleAnnotation(bool enable)
  {
    annotation_enabled_ = enable;
    return enable ? "Annotation enabled" : "Annotation disabled";
  }
  bool annotationEnabled() const { return annotation_enabled_; }

private:
  CloudSink output_;
  ImageSink debug_sink_;
  Detector detector_;
  bool annotation_enabled_;
  pcs_detection::PointCloudAnnotator annotator_;
};
}  // namespace pcs_ros
#endif
