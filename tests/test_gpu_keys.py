"""The voxel key of the CUDA path on every voxel boundary of three resolutions, against exact rational arithmetic (the CPU
twin of this test checks the oracle the same way: tests/test_oracle_keys.py)."""
import numpy as np
import pytest

from util import expected_key_histogram, key_boundary_cloud

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("res", [0.01, 0.005, 0.002])
def test_every_voxel_boundary_against_exact_arithmetic(ctx, pcs, res):
    cloud = key_boundary_cloud(pcs.POINT_DTYPE, res)
    want_keys, want_counts = expected_key_histogram(cloud["x"], res)
    vmap = ctx.create_map(res, 1 << 17)
    vmap.integrate(cloud)  # the stand-alone integrate kernel
    keys, c0, c1, _ = vmap.export()
    order = np.argsort(keys & np.uint64(0xFFFF))
    assert np.array_equal((keys & np.uint64(0xFFFF))[order], want_keys)
    assert np.array_equal((c0.astype(np.uint64) + c1)[order], want_counts)
    # the fused annotate + integrate kernel computes the same keys (the cloud as one organized 3 x n frame)
    vmap.clear()
    n = cloud.size // 3
    ctx.annotate_frames(cloud[: 3 * n], 1, n, 3, None, pcs.annotate_params(), vmap, want_annotated=False, want_mask=False)
    k2, a0, a1, _ = vmap.export()
    o2 = np.argsort(k2 & np.uint64(0xFFFF))
    w2, wc2 = expected_key_histogram(cloud["x"][: 3 * n], res)
    assert np.array_equal((k2 & np.uint64(0xFFFF))[o2], w2) and np.array_equal((a0.astype(np.uint64) + a1)[o2], wc2)
    vmap.destroy()
