"""Python mirrors of the reference's 2D detectors, running on the GPU through the C ABI.

* ``HistogramBackprojection`` keeps the class name, constructor and ``annotate_image`` of
  pcs_detection/src_python/pcs_detection/models/histogram_backprojection.py:4-52, so
  pcs_ros/src/histogram_backprojection_node (``self.histogram_backprojection.annotate_image(np.asarray(cv_image))``,
  :41) works unchanged with this class imported instead.
* ``hsv_thresholding_detector`` is the Python face of pcs_detection::hsvThresholdingDetector
  (pcs_detection/include/pcs_detection/hsv_thresholding.h:45-70) and ``process_image`` of the detector service body
  (pcs_ros/src/hsv_thresholding_node.cpp:8-28).
"""
from __future__ import annotations

import numpy as np

from . import capi

_ctx = None


def default_context() -> capi.Context:
    global _ctx
    if _ctx is None:
        _ctx = capi.Context(0)
    return _ctx


class HistogramBackprojection:
    """Annotates an image based on a previously trained (hue, saturation) histogram."""

    histogram = None
    threshold_min = 30
    threshold_max = 150

    def __init__(self, hist_filepath, context: capi.Context | None = None):
        """Takes an .npy file (or an array) that contains the trained histogram."""
        self.histogram = np.load(hist_filepath) if isinstance(hist_filepath, (str, bytes)) else np.asarray(hist_filepath)
        self.histogram = np.ascontiguousarray(self.histogram, np.float32)
        if self.histogram.ndim != 2:
            raise ValueError("histogram must be 2-D (hue bins x saturation bins)")
        self._ctx = context

    def annotate_image(self, input_image: np.ndarray) -> np.ndarray:
        """Binary 3-channel mask, 255 inside the histogram's colours, same size as the input BGR image."""
        ctx = self._ctx or default_context()
        img = np.ascontiguousarray(input_image, np.uint8)
        return ctx.histogram_backprojection(img, self.histogram, self.threshold_min, self.threshold_max)


def hsv_thresholding_detector(input_image: np.ndarray, inverted: bool = False, context: capi.Context | None = None):
    """8UC3 BGR image -> 8UC1 mask (255 = detected unless inverted)."""
    ctx = context or default_context()
    return ctx.hsv_mask(np.ascontiguousarray(input_image, np.uint8), capi.hsv_params(inverted=int(inverted)))


def process_image(input_image: np.ndarray, context: capi.Context | None = None) -> np.ndarray:
    """Detector service body: the HSV mask replicated to three channels."""
    ctx = context or default_context()
    return ctx.hsv_annotation_image(np.ascontiguousarray(input_image, np.uint8))
