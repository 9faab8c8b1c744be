"""OcTree-compatible export of the voxel map -- .bt writer and colourised leaf-centre cloud."""
import hashlib

import numpy as np
import pytest

from util import random_cloud

pytestmark = pytest.mark.gpu


def pack(kx, ky, kz):
    return (np.asarray(kz, np.uint64) << np.uint64(32)) | (np.asarray(ky, np.uint64) << np.uint64(16)) | np.asarray(kx, np.uint64)


def test_bt_writer_reproduces_reference_fixture_bytes(ctx, pcs, ref_fixtures):
    """The 64 voxels of box_2m.bt at 0.5 m -> the reference's fixture file, byte for byte (SHA-256 of the original)."""
    g = ref_fixtures
    keys = []
    for x, y, z, d in zip(g["bt_kx"], g["bt_ky"], g["bt_kz"], g["bt_depth"]):
        s = 1 << (16 - int(d))
        r = np.arange(s)
        gx, gy, gz = np.meshgrid(int(x) + r, int(y) + r, int(z) + r, indexing="ij")
        keys.append(pack(gx.ravel(), gy.ravel(), gz.ravel()))
    keys = np.concatenate(keys)
    vmap = ctx.create_map(float(g["bt_resolution"]), 1024)
    vmap.merge(keys, None, np.ones(keys.size, np.uint32))
    data = vmap.write_bt()
    assert len(data) == int(g["bt_size"])
    assert hashlib.sha256(data).hexdigest() == str(g["bt_sha256"])
    vmap.destroy()


@pytest.mark.parametrize("seed", [0, 1, 2])
def test_bt_writer_matches_octomap_rules(ctx, pcs, O, seed):
    """Random sparse voxels plus complete 2^3 / 4^3 blocks (which must collapse) vs the oracle's octomap writeBinary."""
    rng = np.random.default_rng(seed)
    k = rng.integers(32768 - 40, 32768 + 40, (3000, 3))
    blocks = []
    for b in range(6):
        base = (rng.integers(32768 - 30, 32768 + 30, 3) // 4) * 4
        s = 4 if b % 2 else 2
        r = np.arange(s)
        gx, gy, gz = np.meshgrid(base[0] + r, base[1] + r, base[2] + r, indexing="ij")
        blocks.append(np.stack([gx.ravel(), gy.ravel(), gz.ravel()], 1))
    k = np.unique(np.concatenate([k] + blocks), axis=0)
    keys = pack(k[:, 0], k[:, 1], k[:, 2])
    vmap = ctx.create_map(0.02, 1 << 14)
    # voxels seen but never labelled (count1 == 0) are not part of the tree
    unl = pack(k[:50, 0] + 100, k[:50, 1], k[:50, 2])
    vmap.merge(np.concatenate([keys, unl]), np.ones(keys.size + 50, np.uint32),
               np.concatenate([np.ones(keys.size, np.uint32), np.zeros(50, np.uint32)]))
    tree = O.Octree(0.02)
    for x, y, z in k:
        tree.insert_key(int(x), int(y), int(z))
    want = tree.write_bt()
    got = vmap.write_bt()
    assert got == want
    again = O.Octree.read_bt(got)
    kk, _ = again.leaf_voxels()
    assert np.array_equal(kk, np.sort(keys))
    vmap.destroy()


def test_empty_map_bt_and_centres(ctx, pcs, O):
    vmap = ctx.create_map(0.1, 64)
    data = vmap.write_bt()
    assert data.endswith(b"size 0\nres 0.1\ndata\n")
    assert vmap.leaf_centers().size == 0
    vmap.destroy()


def test_leaf_centres_round_trip(ctx, pcs, O):
    """centres -> setOctree(centres, res, 255, 255) rebuilds the occupied set; the unlabelled voxels come back as count0."""
    rng = np.random.default_rng(5)
    res = 0.005
    cloud = random_cloud(pcs.POINT_DTYPE, 50000, rng, extent=0.4)
    pose = O.synth_pose(33)
    vmap = ctx.create_map(res, 1 << 17)
    vmap.integrate(cloud, pose, 0, 127)
    keys, c0, c1, _ = vmap.export()
    pts = vmap.leaf_centers()
    assert pts.size == keys.size
    kx = (keys & np.uint64(0xFFFF)).astype(np.int64)
    assert np.array_equal(pts["x"], ((kx - 32768 + 0.5) * res).astype(np.float32))  # octomap keyToCoord
    assert np.array_equal(pts["r"] == 255, c1 > 0) and np.array_equal(pts["r"] == 0, c1 == 0) and (pts["a"] == 255).all()
    back = ctx.create_map(res, 1 << 17)
    back.integrate(pts, None, 255, 255)
    k2, b0, b1, _ = back.export()
    assert np.array_equal(k2, keys)
    assert np.array_equal(b1 > 0, c1 > 0) and np.array_equal(b0 > 0, c1 == 0)
    for m in (vmap, back):
        m.destroy()
