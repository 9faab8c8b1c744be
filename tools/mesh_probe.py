"""Tuning aid: builds a 40-frame map and runs pcs_mesh_contacts_dev on the 2 M-triangle height-field a few times."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import bench
import point_cloud_segmentation_b200 as pcs

ctx = pcs.Context(0)
n, w, h = 40, 1280, 720
poses = np.stack([bench.synth_pose(f) for f in range(n)])
d_pose = ctx.dev_alloc(poses.nbytes); ctx.h2d(d_pose, poses)
d_in = ctx.dev_alloc(n * w * h * 32)
ctx.synth_frames_dev(bench.SEED, 0, n, w, h, d_pose, d_in)
vmap = ctx.create_map(0.005, 1 << 20)
ctx.annotate_frames_dev(d_in, n, w, h, d_pose, pcs.annotate_params(lower=255, upper=255), vmap, None, None)
print(bench.run_mesh(ctx, pcs, vmap, n, 6553.0))
