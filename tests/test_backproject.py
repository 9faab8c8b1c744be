"""histogram back-projection detector.  Oracle pinned on cv2 goldens (CPU); CUDA path vs oracle and
goldens (GPU), bit-exact."""
import os

import numpy as np
import pytest

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "backproject_cv2.npz")


@pytest.fixture(scope="module")
def bp_golden():
    return np.load(GOLDEN)


def cases(g):
    for hi, si in g["cases"]:
        yield g[f"hist{hi}"], g[f"img{hi}_{si}"], g[f"out{hi}_{si}"]
    for hi in sorted(set(int(c[0]) for c in g["cases"])):
        yield g[f"hist{hi}"], g[f"img{hi}_c"], g[f"out{hi}_c"]


def test_oracle_matches_cv2(O, bp_golden):
    n = 0
    for hist, img, want in cases(bp_golden):
        assert np.array_equal(O.histogram_backprojection(img, hist), want)
        n += 1
    assert n == 12
    # an image that is detected everywhere normalises to all zeros (min == max), as cv2 does
    assert bp_golden["out0_c"].max() == 0


@pytest.mark.gpu
def test_gpu_matches_cv2_goldens(ctx, bp_golden):
    for hist, img, want in cases(bp_golden):
        assert np.array_equal(ctx.histogram_backprojection(np.ascontiguousarray(img), hist), want)


@pytest.mark.gpu
@pytest.mark.parametrize("shape", [(1, 1), (2, 9), (5, 7), (33, 129), (480, 640), (721, 1283)])
@pytest.mark.parametrize("bins", [(180, 256), (37, 41)])
def test_gpu_vs_oracle(ctx, O, bp_golden, shape, bins):
    rng = np.random.default_rng(shape[0] * 7 + bins[0])
    h, w = shape
    hist = (rng.random(bins) ** 6 * 300).astype(np.float32)  # mostly small, some > 255 (saturation), .5 ties
    hist[rng.random(bins) < 0.05] = 30.5
    img = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
    img[rng.random((h, w)) < 0.3] = (140, 90, 50)
    big = np.zeros((h, w + 5, 3), np.uint8)
    big[:, 2:2 + w] = img
    view = big[:, 2:2 + w]  # pitched rows
    for tmin, tval in ((30, 150), (0, 255), (200, 1)):
        got = ctx.histogram_backprojection(view, hist, tmin, tval)
        want = O.histogram_backprojection(img, hist, tmin, tval)
        assert np.array_equal(got, want), (tmin, tval)


@pytest.mark.gpu
def test_python_detector_classes_mirror_the_reference(ctx, O, bp_golden, tmp_path):
    from point_cloud_segmentation_b200.detection import HistogramBackprojection, hsv_thresholding_detector, process_image

    path = str(tmp_path / "hist.npy")
    np.save(path, bp_golden["hist1"])
    det = HistogramBackprojection(path, ctx)  # models/histogram_backprojection.py:15-20
    assert (det.threshold_min, det.threshold_max) == (30, 150)
    assert np.array_equal(det.annotate_image(bp_golden["img1_0"]), bp_golden["out1_0"])
    img = bp_golden["img0_2"]
    m = hsv_thresholding_detector(img, context=ctx)
    assert np.array_equal(m, O.hsv_mask(img))
    assert np.array_equal(process_image(img, ctx)[..., 1], m)
