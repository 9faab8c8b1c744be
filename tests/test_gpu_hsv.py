"""Stage 1 parity: pcs_hsv_mask (C ABI, CUDA) vs the oracle and the cv2 golden vectors.  Bit-exact."""
import hashlib

import numpy as np
import pytest

from test_oracle_hsv import all_colours_image

pytestmark = pytest.mark.gpu


def test_all_2_24_colours(ctx, O, hsv_golden):
    img = all_colours_image()  # 4096 x 4096: every BGR colour once
    mask = ctx.hsv_mask(img)
    assert hashlib.sha256(mask.tobytes()).hexdigest() == str(hsv_golden["exhaustive_mask_sha256"])
    assert np.array_equal(mask, O.hsv_mask(img))
    assert int((mask == 255).sum()) == 238377


def test_golden_images_threshold_open_invert(ctx, pcs, hsv_golden):
    for i in range(int(hsv_golden["n_images"])):
        img = np.ascontiguousarray(hsv_golden[f"img{i}"])
        assert np.array_equal(ctx.hsv_mask(img), hsv_golden[f"mask{i}"]), i
        assert np.array_equal(ctx.hsv_mask(img, pcs.hsv_params(morph_open=1)), hsv_golden[f"open{i}"]), i
        assert np.array_equal(ctx.hsv_mask(img, pcs.hsv_params(inverted=1)), 255 - hsv_golden[f"mask{i}"]), i


@pytest.mark.parametrize("shape", [(1, 1), (1, 2), (3, 5), (8, 128), (9, 129), (17, 255), (480, 640), (721, 1283)])
@pytest.mark.parametrize("morph", [0, 1])
def test_ragged_shapes_vs_oracle(ctx, pcs, O, shape, morph):
    rng = np.random.default_rng(shape[0] * 1000 + shape[1])
    h, w = shape
    img = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
    img[rng.random((h, w)) < 0.4] = (140, 90, 50)
    got = ctx.hsv_mask(img, pcs.hsv_params(morph_open=morph, inverted=morph))
    want = O.hsv_mask(img, O.default_params(morph_open=morph, inverted=morph))
    assert np.array_equal(got, want)


def test_row_pitch_and_custom_thresholds(ctx, pcs, O):
    rng = np.random.default_rng(3)
    h, w = 33, 77
    big = rng.integers(0, 256, (h, w + 11, 3), dtype=np.uint8)
    view = big[:, 3:3 + w, :]  # strided rows, unaligned start (cv::Mat ROI)
    p = pcs.hsv_params(h_lo=0, h_hi=60, s_lo=10, s_hi=255, v_lo=0, v_hi=200)
    got = ctx.hsv_mask(view, p)
    want = O.hsv_mask(np.ascontiguousarray(view), O.default_params(h_lo=0, h_hi=60, s_lo=10, s_hi=255, v_lo=0, v_hi=200))
    assert np.array_equal(got, want)
    # strided output keeps the caller's padding bytes
    out = np.full((h, w + 5), 7, np.uint8)
    ctx.hsv_mask_into(view, out[:, :w], p)
    assert np.array_equal(out[:, :w], want) and (out[:, w:] == 7).all()


def test_service_annotation_image_is_mask_x3(ctx, O):
    # pcs_ros/src/hsv_thresholding_node.cpp:20-23
    rng = np.random.default_rng(4)
    img = rng.integers(0, 256, (50, 70, 3), dtype=np.uint8)
    img[10:30, 10:40] = (140, 90, 50)
    ann = ctx.hsv_annotation_image(img)
    m = O.hsv_mask(img)
    assert ann.shape == (50, 70, 3) and all(np.array_equal(ann[..., c], m) for c in range(3))


def test_bad_arguments(ctx, pcs):
    import ctypes as C

    lib = ctx.lib
    p = pcs.hsv_params()
    buf = np.zeros(16, np.uint8)
    vp = buf.ctypes.data_as(C.c_void_p)
    assert lib.pcs_hsv_mask(ctx.h, None, 4, 4, 12, C.byref(p), vp, 4) == pcs.capi.PCS_ERR_BAD_ARG
    assert lib.pcs_hsv_mask(ctx.h, vp, 0, 4, 12, C.byref(p), vp, 4) == pcs.capi.PCS_ERR_BAD_ARG
    assert lib.pcs_hsv_mask(ctx.h, vp, 4, 4, 11, C.byref(p), vp, 4) == pcs.capi.PCS_ERR_BAD_ARG
    assert b"bad argument" in lib.pcs_last_error()


@pytest.mark.parametrize("thr", [
    dict(h_lo=0, h_hi=10, s_lo=0, s_hi=255, v_lo=0, v_hi=255),       # includes hue 0 (both sides of the wrap)
    dict(h_lo=165, h_hi=179, s_lo=30, s_hi=255, v_lo=10, v_hi=250),   # hues that come from negative numerators
    dict(h_lo=0, h_hi=179, s_lo=0, s_hi=255, v_lo=0, v_hi=255),       # everything
    dict(h_lo=100, h_hi=90, s_lo=0, s_hi=255, v_lo=0, v_hi=255),      # empty hue interval
    dict(h_lo=20, h_hi=160, s_lo=200, s_hi=210, v_lo=77, v_hi=77),    # a single value plane
    dict(h_lo=-5, h_hi=300, s_lo=-1, s_hi=999, v_lo=-3, v_hi=256),    # bounds outside the representable ranges
])
def test_decision_tables_all_colours_custom_thresholds(ctx, pcs, O, thr):
    """The contiguous fast path folds the thresholds into decision tables; every colour must still agree with the
    reference arithmetic for arbitrary bounds."""
    img = all_colours_image()
    got = ctx.hsv_mask(img, pcs.hsv_params(**thr))
    want = O.hsv_mask(img, O.default_params(**thr))
    assert np.array_equal(got, want)
    inv = ctx.hsv_mask(img[:100], pcs.hsv_params(inverted=1, **thr))
    assert np.array_equal(inv, 255 - want[:100])


@pytest.mark.parametrize("n_px", [1, 15, 16, 17, 4095, 4096, 4097, 100003])
def test_flat_path_ragged_lengths(ctx, pcs, O, n_px):
    rng = np.random.default_rng(n_px)
    img = rng.integers(0, 256, (1, n_px, 3), dtype=np.uint8)
    img[0, rng.random(n_px) < 0.5] = (140, 90, 50)
    assert np.array_equal(ctx.hsv_mask(img), O.hsv_mask(img))
