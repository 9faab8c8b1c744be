"""The CUDA scan generator (benchmark input) is bit-identical to the oracle's, so the benchmark inputs are the
inputs the parity tests use."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("shape", [(64, 48), (640, 480), (1280, 720)])
def test_frames_bit_identical(ctx, pcs, O, shape):
    w, h = shape
    first, n = 5, 3
    poses = np.stack([O.synth_pose(first + f) for f in range(n)])
    d_pose = ctx.dev_alloc(poses.nbytes)
    d_out = ctx.dev_alloc(n * w * h * 32)
    ctx.h2d(d_pose, poses)
    ctx.synth_frames_dev(O.SEED, first, n, w, h, d_pose, d_out)
    got = np.empty(n * w * h, pcs.POINT_DTYPE)
    ctx.d2h(got, d_out)
    want = np.concatenate([O.synth_frame(first + f, w, h, poses[f]) for f in range(n)])
    assert got.tobytes() == want.tobytes()
    ctx.dev_free(d_pose)
    ctx.dev_free(d_out)
