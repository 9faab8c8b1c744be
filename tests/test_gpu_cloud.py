"""Stage 2 parity through the C ABI: cloudToImage / imageToCloud / colorPassthrough / transform vs the oracle and the
reference's own gtests.  Bit-exact."""
import numpy as np
import pytest

from util import ramp_cloud, random_cloud, same_bytes

pytestmark = pytest.mark.gpu


def test_round_trip_like_utils_unit(ctx, pcs):
    # pcs_detection/test/utils_unit.cpp:38-64
    cloud = ramp_cloud(pcs.POINT_DTYPE)
    pos, bgr = ctx.cloud_to_image(cloud, 255, 1)
    assert pos.shape[:2] == bgr.shape[:2]
    out = ctx.image_to_cloud(bgr, pos)
    assert out.size == cloud.size
    for f in ("x", "y", "z", "r", "g", "b"):
        assert np.array_equal(out[f], cloud[f])


@pytest.mark.parametrize("shape", [(255, 1), (1, 255), (64, 48), (641, 3)])
def test_cloud_image_vs_oracle(ctx, pcs, O, shape):
    w, h = shape
    rng = np.random.default_rng(w * h)
    cloud = random_cloud(pcs.POINT_DTYPE, w * h, rng)
    pos, bgr = ctx.cloud_to_image(cloud, w, h)
    opos, obgr = O.cloud_to_image(cloud, w, h)
    assert same_bytes(pos, opos) and same_bytes(bgr, obgr)
    ann = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
    assert same_bytes(ctx.image_to_cloud(ann, pos), O.image_to_cloud(ann, opos))


def test_unorganized_1x1_cloud_rejected(ctx, pcs):
    # utils.h:103 asserts the cloud is organized
    with pytest.raises(pcs.PcsError) as e:
        ctx.cloud_to_image(np.zeros(1, pcs.POINT_DTYPE), 1, 1)
    assert e.value.status == pcs.capi.PCS_ERR_BAD_ARG


def test_color_passthrough_like_reference_unit(ctx, pcs, O):
    # pcs_scan_integration/test/octomap_mesh_masking_unit.cpp:90-135
    cloud = ramp_cloud(pcs.POINT_DTYPE)
    for lower in range(0, 255, 25):
        for upper in range(lower, 255, 25):
            res = ctx.color_passthrough(cloud, lower, upper, False)
            assert res.size == upper - lower + 1
            assert same_bytes(res, O.color_passthrough(cloud, lower, upper, False))
            neg = ctx.color_passthrough(cloud, lower, upper, True)
            assert neg.size == cloud.size - (upper - lower) + (0 if upper == lower else 1)
            assert same_bytes(neg, O.color_passthrough(cloud, lower, upper, True))


@pytest.mark.parametrize("n", [0, 1, 31, 2048, 2049, 100003])
def test_color_passthrough_order_preserving(ctx, pcs, O, n):
    rng = np.random.default_rng(n + 1)
    cloud = random_cloud(pcs.POINT_DTYPE, n, rng)
    cloud["pad"][:, 0] = np.arange(n)  # payload travels with the point
    got = ctx.color_passthrough(cloud, 40, 210)
    assert same_bytes(got, O.color_passthrough(cloud, 40, 210))
    got = ctx.color_passthrough(cloud, 100, 150, True)
    assert same_bytes(got, O.color_passthrough(cloud, 100, 150, True))


def test_color_passthrough_lower_gt_upper(ctx, pcs):
    # octomap_mesh_masking.cpp:65-69
    with pytest.raises(pcs.PcsError) as e:
        ctx.color_passthrough(ramp_cloud(pcs.POINT_DTYPE), 200, 100)
    assert e.value.status == pcs.capi.PCS_ERR_BAD_ARG and "ower limit greater" in str(e.value)


def test_transform_bit_exact(ctx, pcs, O):
    rng = np.random.default_rng(9)
    cloud = random_cloud(pcs.POINT_DTYPE, 50000, rng, extent=3.0)
    for f in range(4):
        pose = O.synth_pose(f * 37)
        assert same_bytes(ctx.transform_cloud(pose, cloud), O.transform_points(pose, cloud))
    T = rng.normal(size=16).astype(np.float32)
    got, want = ctx.transform_cloud(T, cloud), O.transform_points(T, cloud)
    assert same_bytes(got, want)  # bit-exact, hence also within the 1e-5 m tolerance north_star states
