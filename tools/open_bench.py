"""Tuning aid: device-resident fused annotate+integrate with and without the 5x5 opening (annotate_open_kernel vs
annotate_kernel) on the default scan shape; PCS_OPEN_SECTION_ROWS sweeps the section height."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import point_cloud_segmentation_b200 as pcs
import bench

n_frames, w, h, res = int(os.environ.get("FRAMES", 300)), 1280, 720, 0.005
ctx = pcs.Context(0)
vmap = ctx.create_map(res, 1 << 20)
poses = np.stack([bench.synth_pose(f) for f in range(n_frames)])
n_pts = n_frames * w * h
d_pose, d_in, d_ann, d_mask = ctx.dev_alloc(poses.nbytes), ctx.dev_alloc(n_pts * 32), ctx.dev_alloc(n_pts * 32), ctx.dev_alloc(n_pts)
ctx.h2d(d_pose, poses)
ctx.synth_frames_dev(bench.SEED, 0, n_frames, w, h, d_pose, d_in)
ctx.sync()


def run(label, morph, outputs=True):
    p = pcs.annotate_params(pcs.hsv_params(morph_open=morph), lower=255, upper=255)
    args = (d_in, n_frames, w, h, d_pose, p, vmap, d_ann if outputs else None, d_mask if outputs else None)
    for _ in range(3):
        vmap.clear()
        ctx.annotate_frames_dev(*args)
    evs = []
    for _ in range(10):
        vmap.clear()
        a, b = ctx.event(), ctx.event()
        ctx.record(a)
        ctx.annotate_frames_dev(*args)
        ctx.record(b)
        evs.append((a, b))
    ctx.sync()
    ms = float(np.mean([ctx.elapsed_ms(a, b) for a, b in evs]))
    bpp = 65 if outputs else 32
    print(f"{label}: {ms:.3f} ms  {n_pts / ms / 1e6:.1f} G points/s  {bpp * n_pts / ms / 1e6:.0f} GB/s ({bpp} B/pt)  voxels={vmap.size()}", flush=True)
    return ms


base = run("plain   ", 0)
for rows in os.environ.get("SECTIONS", "0").split(","):
    if int(rows):
        os.environ["PCS_OPEN_SECTION_ROWS"] = rows
    else:
        os.environ.pop("PCS_OPEN_SECTION_ROWS", None)
    ms = run(f"open r={rows:>3}", 1)
    print(f"   ratio to plain: {ms / base:.3f}")
run("plain map-only", 0, outputs=False)
os.environ.pop("PCS_OPEN_SECTION_ROWS", None)
run("open  map-only", 1, outputs=False)
