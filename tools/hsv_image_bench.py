"""Tuning aid: device-resident throughput of the stage-1 image kernel (pcs_hsv_mask_dev), with and without the opening."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import ctypes as C
import numpy as np
import point_cloud_segmentation_b200 as pcs

ctx = pcs.Context(0)
for (w, h, n) in [(640, 480, 64), (1280, 720, 64), (3840, 2160, 16)]:
    img = np.random.default_rng(0).integers(0, 256, (n * h, w, 3), dtype=np.uint8)  # n frames stacked: one tall image
    d_in, d_out = ctx.dev_alloc(img.nbytes), ctx.dev_alloc(n * h * w)
    ctx.h2d(d_in, img)
    for morph in (0, 1):
        p = pcs.hsv_params(morph_open=morph)
        call = lambda: pcs.capi._check(ctx.lib.pcs_hsv_mask_dev(ctx.h, C.c_void_p(d_in), w, n * h, w * 3, C.byref(p), C.c_void_p(d_out), w))
        for _ in range(3):
            call()
        e0, e1 = ctx.event(), ctx.event()
        ctx.record(e0)
        for _ in range(10):
            call()
        ctx.record(e1)
        ms = ctx.elapsed_ms(e0, e1) / 10
        px = n * h * w
        print(f"{w}x{h} x{n} open={morph}: {ms:.3f} ms  {px / ms / 1e6:.1f} G px/s  {4 * px / ms / 1e6:.0f} GB/s")
    ctx.dev_free(d_in); ctx.dev_free(d_out)
