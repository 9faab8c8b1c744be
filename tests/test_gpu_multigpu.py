"""Multi-GPU exchange on real GPUs (needs >= 2 devices; skipped on the single-GPU box): NCCL-record and peer-memory
forms of the owner-partitioned reduction both reproduce the map a single GPU builds from all frames."""
import os
import subprocess
import sys

import numpy as np
import pytest

import bench

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def n_gpus():
    try:
        out = subprocess.run(["nvidia-smi", "-L"], capture_output=True, text=True, timeout=20).stdout
        return sum(1 for l in out.splitlines() if l.startswith("GPU "))
    except Exception:
        return 0


@pytest.mark.skipif(n_gpus() < 2, reason="needs at least 2 GPUs")
def test_exchange_world2_nccl_and_peer_memory(ctx, pcs, tmp_path):
    world = 2
    env = dict(os.environ, OUT_DIR=str(tmp_path))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr",
           "127.0.0.1", "--master-port", "29533", os.path.join(ROOT, "tests", "multigpu_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    # the single-GPU answer over all frames
    w, h, res, per_rank, cap = 320, 240, 0.005, 3, 1 << 18
    n = world * per_rank
    poses = np.stack([bench.synth_pose(f) for f in range(n)])
    d_pose, d_in = ctx.dev_alloc(poses.nbytes), ctx.dev_alloc(n * w * h * 32)
    ctx.h2d(d_pose, poses)
    ctx.synth_frames_dev(bench.SEED, 0, n, w, h, d_pose, d_in)
    ref = ctx.create_map(res, cap)
    ctx.annotate_frames_dev(d_in, n, w, h, d_pose, pcs.annotate_params(lower=255, upper=255), ref, None, None)
    keys, c0, c1, _ = ref.export()
    outs = [np.load(os.path.join(tmp_path, f"rank{r}.npz")) for r in range(world)]
    for r, o in enumerate(outs):
        mine = np.array([pcs.capi.mix64(int(k)) % world == r for k in keys])
        for tag in ("1", "2", "3"):  # NCCL records, peer memory with mailboxes, peer memory with remote atomics
            assert np.array_equal(o["k" + tag], keys[mine]), (r, tag)
            assert np.array_equal(o["a" + tag], c0[mine]) and np.array_equal(o["b" + tag], c1[mine]), (r, tag)
        assert np.array_equal(o["kf"], keys) and np.array_equal(o["af"], c0) and np.array_equal(o["bf"], c1)
    # sharded mesh mask == the one-GPU mask, flag for flag on every rank; rank 0's emitted mesh == the host call's
    from util import grid_mesh, surface

    verts, faces = grid_mesh(81, 41, -0.45, 0.45, -0.25, 0.25, surface)
    faces = np.ascontiguousarray(np.concatenate([faces, faces[:1]]))
    ov, of, _, flags = ctx.mesh_mask(ref, verts, faces, pcs.RETURN_OUTSIDE)
    assert 0 < flags.sum() < flags.size
    for o in outs:
        assert np.array_equal(o["mesh_flags"], flags)
    assert outs[0]["ov"].tobytes() == ov.tobytes() and np.array_equal(outs[0]["of"], of)
    ref.destroy()
    ctx.dev_free(d_pose)
    ctx.dev_free(d_in)
