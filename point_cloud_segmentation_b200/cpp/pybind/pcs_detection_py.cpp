// pcs_detection_py -- pybind11 module for the Python detection scripts (north_star: "pybind only for the Python detection
// scripts").  It binds the three detector entry points of the C ABI that the reference's Python side
// reaches: the HSV-threshold detector (pcs_detection/include/pcs_detection/hsv_thresholding.h:45-70), the detector service
// body (pcs_ros/src/hsv_thresholding_node.cpp:8-28) and HistogramBackprojection.annotate_image
// (pcs_detection/src_python/pcs_detection/models/histogram_backprojection.py:22-52; called from
// pcs_ros/src/histogram_backprojection_node:41).  Images are numpy arrays exactly as cv_bridge hands them over
// (uint8, h x w x 3, BGR).  Everything runs in libpcs_b200 on the GPU; without a CUDA device the calls raise.
#include <pybind11/numpy.h>
#include <pybind11/pybind11.h>

#include <stdexcept>
#include <string>

#include "pcs_b200.h"

namespace py = pybind11;
using image_t = py::array_t<uint8_t, py::array::c_style | py::array::forcecast>;

namespace
{
pcs_ctx* context()
{
  static pcs_ctx* ctx = nullptr;  // one context per process on $PCS_B200_DEVICE (default 0), like the C++ adapters
  if (!ctx)
  {
    const char* dev = std::getenv("PCS_B200_DEVICE");
    if (pcs_ctx_create(dev ? std::atoi(dev) : 0, nullptr, &ctx) != PCS_OK)
      throw std::runtime_error(std::string("libpcs_b200: ") + pcs_last_error());
  }
  return ctx;
}

void check_image(const image_t& img)
{
  if (img.ndim() != 3 || img.shape(2) != 3 || img.shape(0) < 1 || img.shape(1) < 1)
    throw std::invalid_argument("expected a uint8 BGR image of shape (h, w, 3)");
}

void check(int rc)
{
  if (rc != PCS_OK)
    throw std::runtime_error(pcs_last_error());
}

pcs_hsv_params params_of(const py::kwargs& kw)
{
  pcs_hsv_params p;
  pcs_hsv_default_params(&p);
  for (auto item : kw)
  {
    const std::string k = py::cast<std::string>(item.first);
    const int v = py::cast<int>(item.second);
    if (k == "h_lo") p.h_lo = v;
    else if (k == "h_hi") p.h_hi = v;
    else if (k == "s_lo") p.s_lo = v;
    else if (k == "s_hi") p.s_hi = v;
    else if (k == "v_lo") p.v_lo = v;
    else if (k == "v_hi") p.v_hi = v;
    else if (k == "inverted") p.inverted = v;
    else if (k == "morph_open") p.morph_open = v;
    else throw std::invalid_argument("unknown threshold parameter: " + k);
  }
  return p;
}

py::array_t<uint8_t> hsv_thresholding_detector(const image_t& image, const py::kwargs& kw)
{
  check_image(image);
  const pcs_hsv_params p = params_of(kw);
  const int h = (int)image.shape(0), w = (int)image.shape(1);
  py::array_t<uint8_t> mask({ h, w });
  check(pcs_hsv_mask(context(), image.data(), w, h, (size_t)w * 3, &p, mask.mutable_data(), (size_t)w));
  return mask;
}

py::array_t<uint8_t> process_image(const image_t& image, const py::kwargs& kw)
{
  check_image(image);
  const pcs_hsv_params p = params_of(kw);
  const int h = (int)image.shape(0), w = (int)image.shape(1);
  py::array_t<uint8_t> out({ h, w, 3 });
  check(pcs_hsv_annotation_image(context(), image.data(), w, h, (size_t)w * 3, &p, out.mutable_data()));
  return out;
}

py::array_t<uint8_t> histogram_backprojection(const image_t& image,
                                              const py::array_t<float, py::array::c_style | py::array::forcecast>& histogram,
                                              int threshold_min, int threshold_value)
{
  check_image(image);
  if (histogram.ndim() != 2)
    throw std::invalid_argument("histogram must be 2-D (hue bins x saturation bins)");
  const int h = (int)image.shape(0), w = (int)image.shape(1);
  py::array_t<uint8_t> out({ h, w, 3 });
  check(pcs_histogram_backprojection(context(), image.data(), w, h, (size_t)w * 3, histogram.data(), (int)histogram.shape(0),
                                     (int)histogram.shape(1), threshold_min, threshold_value, out.mutable_data()));
  return out;
}
}  // namespace

PYBIND11_MODULE(pcs_detection_py, m)
{
  m.doc() = "GPU detectors of point_cloud_segmentation (libpcs_b200) for the Python detection scripts";
  m.def("version", []() { return pcs_version(); });
  m.def("hsv_thresholding_detector", &hsv_thresholding_detector, py::arg("image"),
        "uint8 BGR image (h, w, 3) -> uint8 mask (h, w), 255 = detected.  Keyword thresholds h_lo .. v_hi, inverted, morph_open.");
  m.def("process_image", &process_image, py::arg("image"), "detector service body: the mask replicated to three channels");
  m.def("histogram_backprojection", &histogram_backprojection, py::arg("image"), py::arg("histogram"),
        py::arg("threshold_min") = 30, py::arg("threshold_value") = 150,
        "HistogramBackprojection.annotate_image: (h, w, 3) BGR + (hue bins, saturation bins) float32 histogram -> (h, w, 3) mask");
}
