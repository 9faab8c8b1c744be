"""Oracle vs the reference's own gtests for cloud<->image utilities and colorPassthrough."""
import numpy as np

from util import ramp_cloud, random_cloud


def test_cloud_image_round_trip_like_utils_unit(O):
    # pcs_detection/test/utils_unit.cpp:38-64
    cloud = ramp_cloud(O.POINT_DTYPE)
    pos, bgr = O.cloud_to_image(cloud, 255, 1)
    assert pos.shape == (1, 255, 3) and bgr.shape == (1, 255, 3)
    out = O.image_to_cloud(bgr, pos)
    assert out.size == cloud.size
    for f in ("x", "y", "z", "r", "g", "b"):
        assert np.array_equal(out[f], cloud[f])
    # Q4: imageToCloud packs rgb into an int32 -> alpha 0 (utils.h:161-162)
    assert (out["a"] == 0).all() and (out["w"] == 1.0).all()


def test_cloud_to_image_layout(O):
    rng = np.random.default_rng(1)
    cloud = random_cloud(O.POINT_DTYPE, 7 * 5, rng)
    pos, bgr = O.cloud_to_image(cloud, 7, 5)
    assert np.array_equal(bgr[..., 0].ravel(), cloud["b"]) and np.array_equal(bgr[..., 2].ravel(), cloud["r"])
    assert np.array_equal(pos[..., 0].ravel().astype(np.float32).view(np.uint32), cloud["x"].view(np.uint32))


def test_color_passthrough_like_reference_unit(O):
    # pcs_scan_integration/test/octomap_mesh_masking_unit.cpp:90-135
    cloud = ramp_cloud(O.POINT_DTYPE)
    for lower in range(0, 255, 25):
        for upper in range(lower, 255, 25):
            res = O.color_passthrough(cloud, lower, upper, False)
            assert res.size == upper - lower + 1
            assert (res["r"] >= lower).all() and (res["r"] <= upper).all()
            neg = O.color_passthrough(cloud, lower, upper, True)
            assert neg.size == cloud.size - (upper - lower) + (0 if upper == lower else 1)
            assert not ((neg["r"] > lower) & (neg["r"] < upper)).any()


def test_color_passthrough_needs_all_three_channels(O):
    c = np.zeros(3, O.POINT_DTYPE)
    c["r"], c["g"], c["b"] = [10, 10, 10], [10, 200, 10], [10, 10, 10]
    assert O.color_passthrough(c, 0, 100).size == 2
    assert np.array_equal(O.color_passthrough(c, 0, 100)["g"], [10, 10])  # order preserved


def test_transform_matches_float32_left_to_right(O):
    rng = np.random.default_rng(2)
    cloud = random_cloud(O.POINT_DTYPE, 1000, rng)
    T = rng.normal(size=(4, 4)).astype(np.float32)
    out = O.transform_points(T, cloud)
    fin = np.isfinite(cloud["x"]) & np.isfinite(cloud["y"]) & np.isfinite(cloud["z"])
    x, y, z = cloud["x"], cloud["y"], cloud["z"]
    with np.errstate(all="ignore"):
        ex = ((T[0, 0] * x + T[0, 1] * y) + T[0, 2] * z) + T[0, 3]
    assert np.array_equal(out["x"][fin].view(np.uint32), ex[fin].astype(np.float32).view(np.uint32))
    # non-finite points are copied through (pcl::transformPointCloud on a non-dense cloud)
    assert out[~fin].tobytes() == cloud[~fin].tobytes()


def test_annotate_frame_is_mask_replicated(O):
    frame = O.synth_frame(0, 64, 48)
    ann, mask = O.annotate_frame(frame, 64, 48)
    assert np.array_equal(ann["r"], mask.ravel()) and np.array_equal(ann["g"], mask.ravel())
    assert np.array_equal(ann["x"].view(np.uint32), frame["x"].view(np.uint32))  # NaNs survive bit-exactly
    assert 0.02 < (mask == 255).mean() < 0.5
