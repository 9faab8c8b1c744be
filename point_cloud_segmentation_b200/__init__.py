"""point_cloud_segmentation_b200 -- B200-native annotate-and-integrate hot path of
swri-robotics/point_cloud_segmentation.

The product is the CUDA library ``libpcs_b200.so`` behind the C ABI of ``include/pcs_b200.h``; this package holds its
sources (``csrc/``), the in-tree build (``_build``), the ctypes binding (``capi``), Python mirrors of the reference's
``pcs_detection`` / ``pcs_scan_integration`` interfaces (``detection``, ``scan_integration``), the multi-GPU map
exchange (``distributed``) and the C++ adapters that keep the reference's signatures (``cpp/``).
"""
from . import capi  # noqa: F401
from .capi import (AnnotateParams, Context, HsvParams, PcsError, POINT_DTYPE, RETURN_COLORIZED, RETURN_INSIDE,  # noqa: F401
                   RETURN_OUTSIDE, VoxelMap, annotate_params, hsv_params, load_library)

__all__ = ["capi", "Context", "VoxelMap", "HsvParams", "AnnotateParams", "PcsError", "POINT_DTYPE", "hsv_params",
           "annotate_params", "load_library", "RETURN_INSIDE", "RETURN_OUTSIDE", "RETURN_COLORIZED"]
