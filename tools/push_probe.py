"""Tuning aid: the exchange kernels on ONE GPU (the "peer" inbox is a local buffer): push-mailbox, merge, clear, for the
48 650-voxel map of the default scan."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import point_cloud_segmentation_b200 as pcs
import bench

n_frames, w, h, res, cap = int(os.environ.get("FRAMES", 300)), 1280, 720, 0.005, 1 << 20
ctx = pcs.Context(0)
local, own = ctx.create_map(res, cap), ctx.create_map(res, cap)
poses = np.stack([bench.synth_pose(f) for f in range(n_frames)])
d_pose, d_in = ctx.dev_alloc(poses.nbytes), ctx.dev_alloc(n_frames * w * h * 32)
ctx.h2d(d_pose, poses)
ctx.synth_frames_dev(bench.SEED, 0, n_frames, w, h, d_pose, d_in)
ctx.annotate_frames_dev(d_in, n_frames, w, h, d_pose, pcs.annotate_params(lower=255, upper=255), local, None, None)
print("local map", local.size())
world = int(os.environ.get("WORLD", 8))
inbox = ctx.dev_alloc(int(ctx.lib.pcs_map_inbox_bytes(cap, world)))
ctx.memset(inbox, 0, 256)
peers = [0] + [inbox] * (world - 1)


def timed(name, fn, reps=20):
    for _ in range(3):
        fn()
    evs = []
    for _ in range(reps):
        own.clear()
        a, b = ctx.event(), ctx.event()
        ctx.record(a)
        fn()
        ctx.record(b)
        evs.append((a, b))
    ctx.sync()
    print(f"{name}: {np.median([ctx.elapsed_ms(a, b) for a, b in evs]) * 1e3:.1f} us", flush=True)


timed("push_mailbox (rank 0 of %d, no clear)" % world, lambda: ctx.push_mailbox(local, own, 0, peers, cap, False))
timed("merge_inbox", lambda: ctx.merge_inbox(own, inbox, 0, world, cap))
timed("push atomics into 1 local owner", lambda: local.push_to_owners_dev([own]))
timed("clear (after push)", lambda: (local.push_to_owners_dev([own]), own.clear()))
timed("empty clear", lambda: own.clear())
