"""Tuning aid: times pcs_annotate_frames_dev with outputs / map switched off one at a time (300 x 1280x720)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import bench
import point_cloud_segmentation_b200 as pcs

n, w, h, res = (16, 3840, 2160, 0.002) if len(sys.argv) > 1 and sys.argv[1] == '4k' else (300, 1280, 720, 0.005)
ctx = pcs.Context(0)
poses = np.stack([bench.synth_pose(f) for f in range(n)])
d_pose = ctx.dev_alloc(poses.nbytes); ctx.h2d(d_pose, poses)
npts = n * w * h
d_in, d_ann, d_mask = ctx.dev_alloc(npts * 32), ctx.dev_alloc(npts * 32), ctx.dev_alloc(npts)
ctx.synth_frames_dev(bench.SEED, 0, n, w, h, d_pose, d_in)
vmap = ctx.create_map(res, 1 << 20)
prm = pcs.annotate_params(lower=255, upper=255)
e0, e1 = ctx.event(), ctx.event()
for name, m, a, k, bpp in [("map+ann+mask", vmap, d_ann, d_mask, 65), ("ann+mask", None, d_ann, d_mask, 65), ("map+ann", vmap, d_ann, None, 64),
                            ("map only", vmap, None, None, 32), ("mask only", None, None, d_mask, 33), ("ann only", None, d_ann, None, 64),
                            ("map+mask", vmap, None, d_mask, 33)]:
    for _ in range(3):
        ctx.annotate_frames_dev(d_in, n, w, h, d_pose, prm, m, a, k)
    ctx.record(e0)
    for _ in range(10):
        ctx.annotate_frames_dev(d_in, n, w, h, d_pose, prm, m, a, k)
    ctx.record(e1)
    ms = ctx.elapsed_ms(e0, e1) / 10
    print(f"{name:14s} {ms:7.3f} ms  {npts / ms / 1e6:7.1f} G pts/s  {bpp * npts / ms / 1e6:7.0f} GB/s")
