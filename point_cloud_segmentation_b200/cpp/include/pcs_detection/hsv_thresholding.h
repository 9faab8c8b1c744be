// pcs_detection/hsv_thresholding.h -- drop-in for the reference header of the same name
// (pcs_detection/include/pcs_detection/hsv_thresholding.h:45-70): same function, same arguments, same defaults;
// the cvtColor + inRange body is replaced by one call into libpcs_b200 (pcs_hsv_mask).
#ifndef PCS_DETECTION_HSV_THRESHOLDING_H
#define PCS_DETECTION_HSV_THRESHOLDING_H

#include "pcs_compat/types.h"

namespace pcs_detection
{
using pcs_compat::Mat;

/**
 * @brief Detects a colour by HSV thresholding and returns an annotation: 255 = detected, 0 = not (unless inverted).
 * @param input_image 8UC3 BGR image
 * @param mask 8UC1 annotation of the same size (allocated here)
 * @param inverted if true, 0 is returned where the colour is detected
 * @return true on success.  The reference always returns true; here a CUDA failure is logged and returns false.
 */
inline bool hsvThresholdingDetector(const Mat& input_image, Mat& mask, bool inverted = false)
{
  pcs_hsv_params p;
  pcs_hsv_default_params(&p);  // hue [95,115], saturation [95,220], value [40,150] (reference :47-52)
  p.inverted = inverted ? 1 : 0;
  if (input_image.empty() || input_image.type() != pcs_compat::CV_8UC3)
  {
    pcs_compat::logError("hsvThresholdingDetector: input must be a non-empty 8UC3 image");
    return false;
  }
  mask.create(input_image.rows, input_image.cols, pcs_compat::CV_8UC1);
  const int rc = pcs_hsv_mask(pcs_compat::default_ctx(), input_image.data, input_image.cols, input_image.rows,
                              input_image.step, &p, mask.data, mask.step);
  if (rc != PCS_OK)
    pcs_compat::logError(pcs_last_error());
  return rc == PCS_OK;
}
}  // namespace pcs_detection
#endif
