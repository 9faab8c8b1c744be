"""Fused annotate(+integrate) parity through the C ABI.

Masks, point labels, voxel keys and per-voxel label counts: bit-exact vs the oracle.  Transformed coordinates:
north_star allows 1e-5 m, the implementation is bit-exact (no FMA contraction on either side).  Log-odds: 1e-5
allowed, bit-exact obtained."""
import numpy as np
import pytest

from util import random_cloud, same_bytes

pytestmark = pytest.mark.gpu


def make_scan(O, n_frames, w, h, first=0):
    frames = np.concatenate([O.synth_frame(first + f, w, h) for f in range(n_frames)])
    poses = np.stack([O.synth_pose(first + f) for f in range(n_frames)])
    return frames, poses


def oracle_scan(O, frames, poses, n_frames, w, h, res, lower, upper, neg, params=None):
    n = w * h
    anns, masks, recs = [], [], {}
    for f in range(n_frames):
        ann, mask = O.annotate_frame(frames[f * n:(f + 1) * n], w, h, params)
        anns.append(ann)
        masks.append(mask)
        k, c0, c1 = O.count_voxels(ann, res, None if poses is None else poses[f], lower, upper, neg)
        for kk, a, b in zip(k.tolist(), c0.tolist(), c1.tolist()):
            r = recs.setdefault(kk, [0, 0])
            r[0] += a
            r[1] += b
    keys = np.array(sorted(recs), np.uint64)
    order = np.argsort(O.morton48(keys), kind="stable")
    keys = keys[order]
    c0 = np.array([recs[int(k)][0] for k in keys], np.uint32)
    c1 = np.array([recs[int(k)][1] for k in keys], np.uint32)
    return np.concatenate(anns), np.stack(masks), keys, c0, c1


@pytest.mark.parametrize("shape,n_frames", [((64, 48), 1), ((133, 71), 3), ((640, 480), 2)])
def test_annotate_integrate_vs_oracle(ctx, pcs, O, shape, n_frames):
    w, h = shape
    res = 0.01
    frames, poses = make_scan(O, n_frames, w, h)
    vmap = ctx.create_map(res, 1 << 18)
    p = pcs.annotate_params(lower=255, upper=255)
    ann, mask = ctx.annotate_frames(frames, n_frames, w, h, poses, p, vmap)
    o_ann, o_mask, o_keys, o_c0, o_c1 = oracle_scan(O, frames, poses, n_frames, w, h, res, 255, 255, False)
    assert np.array_equal(mask, o_mask)          # masks bit-exact
    assert same_bytes(ann, o_ann)                # labelled organized cloud byte-exact (incl. NaNs, alpha 0, padding)
    keys, c0, c1, lo = vmap.export()
    assert np.array_equal(keys, o_keys)          # voxel keys, in octree (Morton) order
    assert np.array_equal(c0, o_c0) and np.array_equal(c1, o_c1)  # per-label hit counts
    lut = np.array([O.logodds_after_hits(c) for c in range(6)], np.float32)
    assert np.array_equal(lo.view(np.uint32), lut[np.minimum(c1, 5)].view(np.uint32))  # log-odds (tolerance 1e-5: exact)
    n_vox, n_occ = vmap.size()
    assert n_vox == keys.size and n_occ == int((c1 > 0).sum())
    finite = np.isfinite(frames["x"]).sum()
    assert int(c0.sum()) + int(c1.sum()) == finite  # every finite in-range point counted exactly once
    vmap.destroy()


def test_occupied_set_equals_reference_octree(ctx, pcs, O):
    """{key : count1 >= 1} and LUT[min(count1,5)] are what octomap's pointer tree holds (octomap_mesh_masking.cpp:105-116)."""
    w, h, n_frames, res = 160, 120, 3, 0.005
    frames, poses = make_scan(O, n_frames, w, h)
    vmap = ctx.create_map(res, 1 << 17)
    ann, _ = ctx.annotate_frames(frames, n_frames, w, h, poses, pcs.annotate_params(lower=255, upper=255), vmap)
    tree = O.Octree(res)
    for f in range(n_frames):
        world = O.transform_points(poses[f], ann[f * w * h:(f + 1) * w * h])
        tree.insert_cloud(O.color_passthrough(world, 255, 255))
    tree.finalize_tesseract()
    keys, c0, c1, lo = vmap.export()
    occ = np.sort(keys[c1 > 0])
    tkeys, tvals = tree.leaf_voxels()
    assert np.array_equal(tkeys, occ)
    vmap.destroy()


def test_world_frame_output_and_identity_pose(ctx, pcs, O):
    w, h = 96, 64
    frames, poses = make_scan(O, 2, w, h, first=40)
    p = pcs.annotate_params(world_frame_output=1)
    ann, mask = ctx.annotate_frames(frames, 2, w, h, poses, p)
    o_ann = np.concatenate([O.transform_points(poses[f], O.annotate_frame(frames[f * w * h:(f + 1) * w * h], w, h)[0])
                            for f in range(2)])
    assert same_bytes(ann, o_ann)
    fin = np.isfinite(o_ann["x"])
    assert np.abs(ann["x"][fin] - o_ann["x"][fin]).max() <= 1e-5  # the tolerance north_star states
    # no pose: exactly imageToCloud
    ann2, _ = ctx.annotate_frames(frames, 2, w, h, None, pcs.annotate_params())
    assert same_bytes(ann2, np.concatenate([O.annotate_frame(frames[f * w * h:(f + 1) * w * h], w, h)[0] for f in range(2)]))


@pytest.mark.parametrize("flt", [(0, 255, False), (255, 255, False), (0, 0, False), (100, 200, True), (0, 255, True)])
def test_label_filters(ctx, pcs, O, flt):
    lower, upper, neg = flt
    w, h, res = 120, 90, 0.02
    frames, poses = make_scan(O, 2, w, h)
    vmap = ctx.create_map(res, 1 << 16)
    ctx.annotate_frames(frames, 2, w, h, poses, pcs.annotate_params(lower=lower, upper=upper, limit_negative=neg), vmap,
                        want_annotated=False, want_mask=False)
    _, _, o_keys, o_c0, o_c1 = oracle_scan(O, frames, poses, 2, w, h, res, lower, upper, neg)
    keys, c0, c1, _ = vmap.export()
    assert np.array_equal(keys, o_keys) and np.array_equal(c0, o_c0) and np.array_equal(c1, o_c1)
    vmap.destroy()


def test_detector_options_inverted_and_opening(ctx, pcs, O):
    w, h, res = 150, 100, 0.01
    frames, poses = make_scan(O, 2, w, h)
    for inverted, morph in ((1, 0), (0, 1), (1, 1)):
        vmap = ctx.create_map(res, 1 << 16)
        p = pcs.annotate_params(pcs.hsv_params(inverted=inverted, morph_open=morph), lower=255, upper=255)
        ann, mask = ctx.annotate_frames(frames, 2, w, h, poses, p, vmap)
        op = O.default_params(inverted=inverted, morph_open=morph)
        o_ann, o_mask, o_keys, o_c0, o_c1 = oracle_scan(O, frames, poses, 2, w, h, res, 255, 255, False, op)
        assert np.array_equal(mask, o_mask) and same_bytes(ann, o_ann)
        keys, c0, c1, _ = vmap.export()
        assert np.array_equal(keys, o_keys) and np.array_equal(c1, o_c1) and np.array_equal(c0, o_c0)
        vmap.destroy()


@pytest.mark.parametrize("w,h,n_frames,section_rows", [(700, 420, 2, 0), (300, 200, 3, 180), (513, 97, 1, 0), (256, 33, 2, 12),
                                                        (260, 64, 1, 64), (1283, 50, 1, 0)])
def test_fused_opening_bands_and_sections(ctx, pcs, O, monkeypatch, w, h, n_frames, section_rows):
    """The opening needs a 9x9 neighbourhood: bands of 256 columns and row sections must stitch without seams, at frame
    borders and at every band / section boundary, with all outputs (mask, annotated cloud, voxel counts)."""
    if section_rows:
        monkeypatch.setenv("PCS_OPEN_SECTION_ROWS", str(section_rows))
    res = 0.01
    frames, poses = make_scan(O, n_frames, w, h)
    rng = np.random.default_rng(w * h)
    # blobs and speckles around the threshold colour so that the opening removes and keeps things everywhere
    hit = rng.random(frames.size) < 0.55
    frames["b"][hit], frames["g"][hit], frames["r"][hit] = 140, 90, 50
    for inverted in (0, 1):
        vmap = ctx.create_map(res, 1 << 17)
        p = pcs.annotate_params(pcs.hsv_params(inverted=inverted, morph_open=1), lower=255, upper=255)
        ann, mask = ctx.annotate_frames(frames, n_frames, w, h, poses, p, vmap)
        op = O.default_params(inverted=inverted, morph_open=1)
        o_ann, o_mask, o_keys, o_c0, o_c1 = oracle_scan(O, frames, poses, n_frames, w, h, res, 255, 255, False, op)
        assert np.array_equal(mask, o_mask)
        assert 0.02 < (o_mask == 255).mean() < 0.98
        assert same_bytes(ann, o_ann)
        keys, c0, c1, _ = vmap.export()
        assert np.array_equal(keys, o_keys) and np.array_equal(c1, o_c1) and np.array_equal(c0, o_c0)
        # mask-only and map-only variants of the kernel give the same answers
        _, mask2 = ctx.annotate_frames(frames, n_frames, w, h, poses, p, None, want_annotated=False)
        assert np.array_equal(mask2, o_mask)
        vmap.clear()
        ctx.annotate_frames(frames, n_frames, w, h, poses, p, vmap, want_annotated=False, want_mask=False)
        k2, a0, a1, _ = vmap.export()
        assert np.array_equal(k2, o_keys) and np.array_equal(a1, o_c1) and np.array_equal(a0, o_c0)
        vmap.destroy()


def test_nan_inf_and_out_of_range_points_are_culled(ctx, pcs, O):
    rng = np.random.default_rng(11)
    w, h = 200, 50
    cloud = random_cloud(pcs.POINT_DTYPE, w * h, rng, extent=400.0, nan_frac=0.2, inf_frac=0.1)  # map spans +-327.68 m
    vmap = ctx.create_map(0.01, 1 << 15)
    ctx.annotate_frames(cloud, 1, w, h, None, pcs.annotate_params(), vmap, want_annotated=False, want_mask=False)
    ann, _ = O.annotate_frame(cloud, w, h)
    o_k, o_c0, o_c1 = O.count_voxels(ann, 0.01)
    keys, c0, c1, _ = vmap.export()
    assert np.array_equal(keys, o_k) and np.array_equal(c0, o_c0) and np.array_equal(c1, o_c1)
    assert 0 < keys.size < w * h * 0.7
    vmap.destroy()


def test_empty_batch_and_bad_args(ctx, pcs):
    ann, mask = ctx.annotate_frames(np.zeros(0, pcs.POINT_DTYPE), 0, 4, 4)
    assert ann.size == 0
    with pytest.raises(pcs.PcsError) as e:
        ctx.annotate_frames(np.zeros(16, pcs.POINT_DTYPE), 1, 4, 4, None, pcs.annotate_params(lower=9, upper=3))
    assert e.value.status == pcs.capi.PCS_ERR_BAD_ARG


def test_accumulation_is_order_and_batching_independent(ctx, pcs, O):
    """setOctree on the concatenation of frames == any batching / ordering of per-frame calls (hits commute)."""
    w, h, n, res = 128, 96, 6, 0.005
    frames, poses = make_scan(O, n, w, h)
    p = pcs.annotate_params(lower=255, upper=255)
    a = ctx.create_map(res, 1 << 17)
    ctx.annotate_frames(frames, n, w, h, poses, p, a, want_annotated=False, want_mask=False)
    b = ctx.create_map(res, 1 << 17)
    for f in reversed(range(n)):
        ctx.annotate_frames(frames[f * w * h:(f + 1) * w * h], 1, w, h, poses[f:f + 1], p, b, want_annotated=False,
                            want_mask=False)
    ea, eb = a.export(), b.export()
    assert all(np.array_equal(x, y) for x, y in zip(ea, eb))
    a.clear()
    assert a.size() == (0, 0) and a.export()[0].size == 0
    a.destroy()
    b.destroy()


@pytest.mark.parametrize("flt", [(255, 255, False), (0, 0, False), (0, 255, False), (10, 200, True)])
def test_labelled_points_compaction(ctx, pcs, O, flt):
    """Stage 2's compacted product: colorPassthrough(transform(annotated frame)) without the points octomap skips, in
    input order, byte for byte."""
    import ctypes as C

    lower, upper, neg = flt
    w, h, n_frames, res = 150, 90, 3, 0.005
    frames, poses = make_scan(O, n_frames, w, h, first=7)
    frames["x"][::97] = 400.0  # outside the +-163.84 m key range at 5 mm
    frames["y"][5::131] = np.inf
    want = []
    inv = 1.0 / res
    for f in range(n_frames):
        ann, _ = O.annotate_frame(frames[f * w * h:(f + 1) * w * h], w, h)
        world = O.transform_points(poses[f], ann)
        kept = O.color_passthrough(world, lower, upper, neg)
        k = C.c_uint16(0)
        ok = np.array([all(O.lib().orc_coord_to_key(inv, float(p[a]), C.byref(k)) for a in ("x", "y", "z")) for p in kept], bool)
        want.append(kept[ok] if kept.size else kept)
    want = np.concatenate(want)
    got = ctx.labelled_points(frames, n_frames, w, h, poses, pcs.annotate_params(lower=lower, upper=upper, limit_negative=neg), res)
    assert same_bytes(got, want)
    assert 0 < got.size < frames.size
    # without the range test only non-finite points are culled
    got2 = ctx.labelled_points(frames, n_frames, w, h, poses, pcs.annotate_params(lower=lower, upper=upper, limit_negative=neg), 0.0)
    assert got2.size >= got.size and np.isfinite(got2["x"]).all()
