"""Pins the oracle's detector restatement (oracle/hsv.c) to cv2's output (tests/golden/hsv_cv2.npz).

hsvThresholdingDetector itself is unpinned by the reference's tests (no test touches it); the golden vectors come
from the OpenCV calls it makes (pcs_detection/include/pcs_detection/hsv_thresholding.h:55-59)."""
import hashlib

import numpy as np


def all_colours_image():
    r = np.arange(256, dtype=np.uint8)
    return np.stack([np.broadcast_to(r[None, None, :], (256, 256, 256)),
                     np.broadcast_to(r[None, :, None], (256, 256, 256)),
                     np.broadcast_to(r[:, None, None], (256, 256, 256))], axis=-1).reshape(4096, 4096, 3).copy()


def test_exhaustive_hsv_and_mask_match_cv2(O, hsv_golden):
    img = all_colours_image()
    hsv = O.bgr2hsv(img)
    assert hashlib.sha256(hsv.tobytes()).hexdigest() == str(hsv_golden["exhaustive_hsv_sha256"])
    mask = O.hsv_mask(img)
    assert hashlib.sha256(mask.tobytes()).hexdigest() == str(hsv_golden["exhaustive_mask_sha256"])
    assert int((mask == 255).sum()) == int(hsv_golden["exhaustive_positive"]) == 238377
    packed = np.packbits(mask.reshape(-1) != 0)
    assert np.array_equal(packed, hsv_golden["exhaustive_mask_packed"])
    h, n = O.hsv_exhaustive_hash()
    assert n == 238377


def test_reference_thresholds_are_the_defaults(O):
    p = O.default_params()
    # hsv_thresholding.h:47-52
    assert (p.h_lo, p.h_hi, p.s_lo, p.s_hi, p.v_lo, p.v_hi) == (95, 115, 95, 220, 40, 150)
    assert p.inverted == 0 and p.morph_open == 0


def test_random_images_and_opening_match_cv2(O, hsv_golden):
    for i in range(int(hsv_golden["n_images"])):
        img = hsv_golden[f"img{i}"]
        assert np.array_equal(O.hsv_mask(img), hsv_golden[f"mask{i}"])
        assert np.array_equal(O.hsv_mask(img, O.default_params(morph_open=1)), hsv_golden[f"open{i}"])
        inv = O.hsv_mask(img, O.default_params(inverted=1))
        assert np.array_equal(inv, 255 - hsv_golden[f"mask{i}"])  # hsv_thresholding.h:61-62


def test_synthetic_paint_colour_is_detected(O):
    # the painted stripes of the synthetic scan must fall inside the reference thresholds for every noise value
    for db in (-8, 0, 8):
        for dg in (-8, 0, 8):
            for dr in (-8, 0, 8):
                px = np.array([[[140 + db, 90 + dg, 50 + dr]]], np.uint8)
                assert O.hsv_mask(px)[0, 0] == 255
