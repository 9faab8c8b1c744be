"""The pybind11 module for the Python detection scripts (pcs_detection_py): same answers as the ctypes binding and the cv2
goldens on the GPU; on a box without a GPU it imports and refuses to compute."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def mod():
    from point_cloud_segmentation_b200 import _build
    from point_cloud_segmentation_b200.cpp import build_cpp

    _build.build()
    build_cpp.build_pybind()
    sys.path.insert(0, os.path.join(ROOT, "point_cloud_segmentation_b200"))
    import pcs_detection_py

    return pcs_detection_py


def test_module_imports_and_has_no_cpu_fallback(mod):
    assert mod.version() == 100
    with pytest.raises(ValueError):
        mod.hsv_thresholding_detector(np.zeros((4, 4), np.uint8))  # not an image
    try:
        out = mod.hsv_thresholding_detector(np.zeros((4, 4, 3), np.uint8))
        assert out.shape == (4, 4) and not out.any()  # running on the GPU box
    except RuntimeError as e:
        assert "no CPU fallback" in str(e)


@pytest.mark.gpu
def test_pybind_equals_ctypes_and_goldens(mod, ctx, pcs, hsv_golden):
    for i in range(int(hsv_golden["n_images"])):
        img = np.ascontiguousarray(hsv_golden[f"img{i}"])
        assert np.array_equal(mod.hsv_thresholding_detector(img), hsv_golden[f"mask{i}"])
        assert np.array_equal(mod.hsv_thresholding_detector(img, morph_open=1), hsv_golden[f"open{i}"])
        assert np.array_equal(mod.hsv_thresholding_detector(img, inverted=1), 255 - hsv_golden[f"mask{i}"])
        assert np.array_equal(mod.process_image(img), ctx.hsv_annotation_image(img))
    g = np.load(os.path.join(ROOT, "tests", "golden", "backproject_cv2.npz"))
    n = 0
    for k in g.files:
        if not k.startswith("img"):
            continue
        h_idx, tag = k[3:].split("_")
        img, hist = np.ascontiguousarray(g[k]), g["hist" + h_idx]
        got = mod.histogram_backprojection(img, hist, 30, 150)
        assert np.array_equal(got, g[f"out{h_idx}_{tag}"]), k  # cv2 golden
        assert np.array_equal(got, ctx.histogram_backprojection(img, hist, 30, 150)), k
        n += 1
    assert n >= 12
