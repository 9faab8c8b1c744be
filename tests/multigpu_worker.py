"""Worker of tests/test_gpu_multigpu.py: run under torch.distributed.run with one process per GPU.

Every rank integrates its own frames, then the owner-partitioned exchange runs twice -- through NCCL records
(distributed.MapExchange) and through peer memory (distributed.PeerMapExchange) -- and both owned slices are written
to OUT_DIR for the parent test to compare with the single-GPU answer."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import bench  # noqa: E402
import point_cloud_segmentation_b200 as pcs  # noqa: E402
from point_cloud_segmentation_b200.distributed import MapExchange, PeerMapExchange  # noqa: E402


def main():
    rank, world, local_rank = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    out_dir = os.environ["OUT_DIR"]
    torch.cuda.set_device(local_rank)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    shared_stream = torch.cuda.Stream()  # libpcs_b200 kernels and NCCL collectives are ordered on this one stream
    torch.cuda.set_stream(shared_stream)
    ctx = pcs.Context(local_rank, shared_stream.cuda_stream)
    w, h, res, per_rank, cap = 320, 240, 0.005, 3, 1 << 18
    first = rank * per_rank
    poses = np.stack([bench.synth_pose(first + f) for f in range(per_rank)])
    d_pose, d_in = ctx.dev_alloc(poses.nbytes), ctx.dev_alloc(per_rank * w * h * 32)
    ctx.h2d(d_pose, poses)
    ctx.synth_frames_dev(bench.SEED, first, per_rank, w, h, d_pose, d_in)
    local = ctx.create_map(res, cap)
    ctx.annotate_frames_dev(d_in, per_rank, w, h, d_pose, pcs.annotate_params(lower=255, upper=255), local, None, None)

    nccl = MapExchange(ctx, local, dist, torch, cap)
    nccl.exchange()
    nccl.exchange()  # repeatable
    k1, a1, b1, _ = nccl._owned_map.export()
    full = ctx.create_map(res, cap)
    from point_cloud_segmentation_b200.distributed import GpuMapAdapter

    nccl.allgather_owned(GpuMapAdapter(ctx, full))
    kf, af, bf, _ = full.export()

    at = PeerMapExchange(ctx, local, dist, torch, cap, mode="atomics", barrier="flags")  # remote atomics, stream-memop barrier
    at.exchange()
    at.exchange()
    ctx.sync()
    k3, a3, b3, _ = at.owned_map.export()
    at.close()
    peer = PeerMapExchange(ctx, local, dist, torch, cap)  # mailbox records, kernel barrier (defaults)
    peer.exchange()
    peer.exchange()
    ctx.sync()
    k2, a2, b2, _ = peer.owned_map.export()
    # sharded mesh mask on the gathered map (peer slices gathered this time): flags + rank 0's emitted OUTSIDE mesh
    from point_cloud_segmentation_b200.distributed import allgather_map, sharded_mesh_flags

    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from util import grid_mesh, surface

    full2 = ctx.create_map(res, cap)
    allgather_map(peer.owned, GpuMapAdapter(ctx, full2), dist, torch, torch.device("cuda", local_rank))
    verts, faces = grid_mesh(81, 41, -0.45, 0.45, -0.25, 0.25, surface)
    faces = np.ascontiguousarray(np.concatenate([faces, faces[:1]]))  # 6643 triangles: not a multiple of 2
    nv, nt = verts.shape[0], faces.shape[0]
    d_v, d_f = ctx.dev_alloc(verts.nbytes), ctx.dev_alloc(faces.nbytes)
    ctx.h2d(d_v, verts)
    ctx.h2d(d_f, faces)
    fl = sharded_mesh_flags(ctx, full2, dist, torch, d_v, nv, d_f, nt)
    ctx.sync()
    mesh_flags = fl.cpu().numpy()
    emitted = {}
    if rank == 0:
        d_ov, d_of, d_rgb, d_n = ctx.dev_alloc(nt * 72), ctx.dev_alloc(nt * 16), ctx.dev_alloc(nt * 9), ctx.dev_alloc(8)
        ctx.mesh_emit_dev(d_v, nv, d_f, nt, fl.data_ptr(), pcs.RETURN_OUTSIDE, nt, d_ov, d_of, d_rgb, d_n)
        n_out = np.zeros(1, np.uint64)
        ctx.d2h(n_out, d_n)
        k = int(n_out[0])
        ov, of = np.empty((3 * k, 3), np.float64), np.empty((k, 4), np.int32)
        ctx.d2h(ov, d_ov)
        ctx.d2h(of, d_of)
        emitted = dict(ov=ov, of=of)
    np.savez(os.path.join(out_dir, f"rank{rank}.npz"), k1=k1, a1=a1, b1=b1, k2=k2, a2=a2, b2=b2, kf=kf, af=af, bf=bf,
             k3=k3, a3=a3, b3=b3, mesh_flags=mesh_flags, **emitted)
    peer.close()
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
