"""Host-side logic of the multi-GPU map exchange on CPU: world_size 2, gloo backend, maps replaced by a numpy
stand-in with the same interface (the kernels themselves are covered by the -m gpu tests)."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class NumpyMap:
    """CPU stand-in for capi.VoxelMap behind distributed.MapExchange (records = {key48, count0 | count1 << 32})."""

    def __init__(self):
        self.recs = {}

    def size(self):
        return len(self.recs)

    def clear(self):
        self.recs = {}

    def add(self, keys, c0, c1):
        for k, a, b in zip(keys.tolist(), c0.tolist(), c1.tolist()):
            self.recs[k] = self.recs.get(k, 0) + (a | (b << 32))

    def pack_by_owner_into(self, n_owners, buf):
        from point_cloud_segmentation_b200 import capi

        keys = np.array(sorted(self.recs), np.uint64)
        owners = np.array([capi.mix64(int(k)) % n_owners for k in keys], np.int64)
        order = np.argsort(owners, kind="stable")
        view = buf.numpy().view(np.uint64)
        view[0:2 * keys.size:2] = keys[order]
        view[1:2 * keys.size:2] = np.array([self.recs[int(k)] for k in keys[order]], np.uint64)
        return np.bincount(owners, minlength=n_owners).astype(np.uint64)

    def merge_records_from(self, buf, n):
        view = buf.numpy().view(np.uint64)
        for i in range(n):
            k, c = int(view[2 * i]), int(view[2 * i + 1])
            self.recs[k] = self.recs.get(k, 0) + c


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch
    import torch.distributed as dist

    from oracle import oracle as O
    from point_cloud_segmentation_b200 import capi
    from point_cloud_segmentation_b200.distributed import MapExchange

    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    w, h, res, per_rank = 96, 64, 0.01, 2
    local = NumpyMap()
    for f in range(rank * per_rank, (rank + 1) * per_rank):  # frames sharded by rank
        ann, _ = O.annotate_frame(O.synth_frame(f, w, h), w, h)
        local.add(*O.count_voxels(ann, res, O.synth_pose(f), 255, 255))
    owned, full = NumpyMap(), NumpyMap()
    ex = MapExchange(None, None, dist, torch, 0, local=local, owned=owned, device=torch.device("cpu"))
    n_recv = ex.exchange()
    assert n_recv >= owned.size() > 0
    assert all(capi.mix64(k) % world == rank for k in owned.recs)  # spatial-hash ownership
    ex.exchange()  # idempotent with clear_owned
    total = ex.allgather_owned(full)
    keys = np.array(sorted(full.recs), np.uint64)
    # sharded mesh mask on the gathered map: each rank decides its range of triangles (the oracle stands in for the
    # kernel), the flags are gathered; nt = 101 is not a multiple of the world size
    from point_cloud_segmentation_b200.distributed import gather_triangle_flags, triangle_range
    from util import grid_mesh, surface

    verts, faces = grid_mesh(10, 5, -0.45, 0.45, -0.25, 0.25, surface)
    faces = np.concatenate([faces, faces[:1]])
    occ = np.array([k for k in keys if (full.recs[int(k)] >> 32) > 0], np.uint64)

    def compute(lo, hi, out):
        out[: hi - lo] = torch.from_numpy(O.mesh_contacts_keys(occ, res, verts, faces[lo:hi]))

    flags = gather_triangle_flags(dist, torch, torch.device("cpu"), faces.shape[0], compute).numpy()
    assert triangle_range(101, world, 0) == (0, 50) and triangle_range(101, world, 1) == (50, 101)
    np.savez(os.path.join(out_dir, f"rank{rank}.npz"), keys=keys,
             counts=np.array([full.recs[int(k)] for k in keys], np.uint64), owned=np.array(sorted(owned.recs), np.uint64),
             total=total, flags=flags)
    dist.barrier()
    dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_exchange_world2_gloo(O, tmp_path):
    import torch.multiprocessing as mp

    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    # single-process answer over all frames
    w, h, res = 96, 64, 0.01
    ref = NumpyMap()
    for f in range(4):
        ann, _ = O.annotate_frame(O.synth_frame(f, w, h), w, h)
        ref.add(*O.count_voxels(ann, res, O.synth_pose(f), 255, 255))
    keys = np.array(sorted(ref.recs), np.uint64)
    counts = np.array([ref.recs[int(k)] for k in keys], np.uint64)
    outs = [np.load(os.path.join(tmp_path, f"rank{r}.npz")) for r in range(world)]
    for o in outs:  # after the all-gather every rank holds the map a single GPU would have built
        assert np.array_equal(o["keys"], keys) and np.array_equal(o["counts"], counts)
        assert int(o["total"]) == keys.size
    # the sharded mask equals the one-process mask, flag for flag, on every rank
    from util import grid_mesh, surface

    verts, faces = grid_mesh(10, 5, -0.45, 0.45, -0.25, 0.25, surface)
    faces = np.concatenate([faces, faces[:1]])
    want = O.mesh_contacts_keys(keys[(counts >> np.uint64(32)) > 0], res, verts, faces)
    assert 0 < want.sum() < want.size
    for o in outs:
        assert np.array_equal(o["flags"], want)
    owned = np.concatenate([o["owned"] for o in outs])
    assert owned.size == keys.size and np.array_equal(np.sort(owned), keys)  # slices partition the map
