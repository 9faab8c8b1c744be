"""Tuning aid: the mesh-mask section of bench.py on a smaller scan (FRAMES, default 100)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import point_cloud_segmentation_b200 as pcs
import bench

n_frames, w, h, res = int(os.environ.get("FRAMES", 100)), 1280, 720, 0.005
ctx = pcs.Context(0)
vmap = ctx.create_map(res, 1 << 20)
poses = np.stack([bench.synth_pose(f) for f in range(n_frames)])
d_pose, d_in = ctx.dev_alloc(poses.nbytes), ctx.dev_alloc(n_frames * w * h * 32)
ctx.h2d(d_pose, poses)
ctx.synth_frames_dev(bench.SEED, 0, n_frames, w, h, d_pose, d_in)
ctx.annotate_frames_dev(d_in, n_frames, w, h, d_pose, pcs.annotate_params(lower=255, upper=255), vmap, None, None)
print("map", vmap.size())
r = bench.run_mesh(ctx, pcs, vmap, n_frames, bench.peak_hbm_gbs()[0], e2e=os.environ.get("E2E", "1") == "1")
for k, c in r["cases"].items():
    print(k, f"inside={c['inside_fraction']:.3f}", f"flags {c['flags']['ms'] * 1e3:.1f} us ({c['flags']['launches']} launches, frac {c['flags']['roofline']['frac']:.2f})",
          f"mask {c['mask']['ms'] * 1e3:.1f} us ({c['mask']['launches']} launches, frac {c['mask']['roofline']['frac']:.2f})",
          f"e2e {c.get('e2e', {}).get('ms', 0):.1f} ms")
