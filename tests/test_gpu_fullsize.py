"""BASELINE.json full-size runs (configs[1]: 300 x 1280x720 at 5 mm; configs[4]-shaped: 3840x2160 at 2 mm) through the
device-pointer C ABI, checked by size-independent properties plus the oracle on sampled frames."""
import numpy as np
import pytest

import bench

pytestmark = pytest.mark.gpu


class Scan:
    def __init__(self, ctx, pcs, n_frames, w, h, first=0):
        self.ctx, self.n, self.w, self.h = ctx, n_frames, w, h
        self.ppf = w * h
        self.poses = np.stack([bench.synth_pose(first + f) for f in range(n_frames)])
        self.d_pose = ctx.dev_alloc(self.poses.nbytes)
        self.d_in = ctx.dev_alloc(n_frames * self.ppf * 32)
        ctx.h2d(self.d_pose, self.poses)
        ctx.synth_frames_dev(bench.SEED, first, n_frames, w, h, self.d_pose, self.d_in)

    def frame(self, pcs, f, d_buf=None):
        out = np.empty(self.ppf, pcs.POINT_DTYPE)
        self.ctx.d2h(out, (d_buf or self.d_in) + f * self.ppf * 32)
        return out

    def free(self):
        self.ctx.dev_free(self.d_pose)
        self.ctx.dev_free(self.d_in)


def export_tuple(m):
    return m.export()


@pytest.mark.parametrize("n_frames,w,h,res,cap", [(300, 1280, 720, 0.005, 1 << 21), (8, 3840, 2160, 0.002, 1 << 22)])
def test_full_size_properties(ctx, pcs, O, n_frames, w, h, res, cap):
    scan = Scan(ctx, pcs, n_frames, w, h)
    n_pts = n_frames * scan.ppf
    d_ann = ctx.dev_alloc(n_pts * 32)
    d_mask = ctx.dev_alloc(n_pts)
    params = pcs.annotate_params(lower=255, upper=255)
    full = ctx.create_map(res, cap)
    ctx.annotate_frames_dev(scan.d_in, n_frames, w, h, scan.d_pose, params, full, d_ann, d_mask)
    keys, c0, c1, lo = export_tuple(full)

    # (1) every finite point is counted exactly once (the synthetic scan stays inside the +-163 m / +-65 m key range)
    xs = np.empty(n_pts * 8, np.float32)
    ctx.d2h(xs, scan.d_in)
    finite = int(np.isfinite(xs[0::8]).sum())
    assert int(c0.sum(dtype=np.uint64)) + int(c1.sum(dtype=np.uint64)) == finite
    assert 0.94 < finite / n_pts < 0.96  # 5 % NaN holes
    del xs

    # (2) labels: count1 == number of detected finite points == number of 255s in the mask at finite points
    mask = np.empty(n_pts, np.uint8)
    ctx.d2h(mask, d_mask)
    zs = np.empty(n_pts * 8, np.float32)
    ctx.d2h(zs, d_ann)
    fin = np.isfinite(zs[0::8])
    assert int(c1.sum(dtype=np.uint64)) == int(((mask == 255) & fin).sum())
    assert set(np.unique(mask).tolist()) <= {0, 255}
    rgba = zs.view(np.uint32)[4::8]
    assert np.array_equal(rgba == 0x00FFFFFF, mask == 255) and np.array_equal(rgba == 0, mask == 0)  # colour := annotation
    del zs, rgba, fin

    # (3) export is in strict Morton order (sorted, unique keys), log-odds is the LUT of count1
    m = O.morton48(keys)
    assert (m[1:] > m[:-1]).all()
    lut = np.array([O.logodds_after_hits(c) for c in range(6)], np.float32)
    assert np.array_equal(lo.view(np.uint32), lut[np.minimum(c1, 5)].view(np.uint32))

    # (4) linearity / order independence: two halves in separate maps, merged == the whole scan; repeat == same
    half = n_frames // 2
    a, b = ctx.create_map(res, cap), ctx.create_map(res, cap)
    ctx.annotate_frames_dev(scan.d_in + half * scan.ppf * 32, n_frames - half, w, h, scan.d_pose + half * 64, params, a, None, None)
    ctx.annotate_frames_dev(scan.d_in, half, w, h, scan.d_pose, params, b, None, None)
    ka, a0, a1, _ = a.export()
    b.merge(ka, a0, a1)
    assert all(np.array_equal(x, y) for x, y in zip(b.export(), (keys, c0, c1, lo)))
    full.clear()
    ctx.annotate_frames_dev(scan.d_in, n_frames, w, h, scan.d_pose, params, full, d_ann, d_mask)
    assert all(np.array_equal(x, y) for x, y in zip(full.export(), (keys, c0, c1, lo)))

    # (5) sampled frames against the oracle, byte for byte
    for f in (0, n_frames // 2, n_frames - 1):
        frame = scan.frame(pcs, f)
        assert frame.tobytes() == O.synth_frame(f, w, h, scan.poses[f]).tobytes()
        o_ann, o_mask = O.annotate_frame(frame, w, h)
        assert scan.frame(pcs, f, d_ann).tobytes() == o_ann.tobytes()
        assert np.array_equal(mask[f * scan.ppf:(f + 1) * scan.ppf].reshape(h, w), o_mask)

    # (6) per-frame map of one sampled frame against the oracle's exact counts
    one = ctx.create_map(res, cap)
    f = n_frames - 1
    ctx.annotate_frames_dev(scan.d_in + f * scan.ppf * 32, 1, w, h, scan.d_pose + f * 64, params, one, None, None)
    o_ann, _ = O.annotate_frame(scan.frame(pcs, f), w, h)
    ok, o0, o1 = O.count_voxels(o_ann, res, scan.poses[f], 255, 255)
    gk, g0, g1, _ = one.export()
    assert np.array_equal(gk, ok) and np.array_equal(g0, o0) and np.array_equal(g1, o1)

    for mm in (full, a, b, one):
        mm.destroy()
    ctx.dev_free(d_ann)
    ctx.dev_free(d_mask)
    scan.free()


def test_full_size_scan_with_the_opening(ctx, pcs, O):
    """configs[1] at full size through annotate_open_kernel (one launch for 300 frames, bands x sections x frames CTAs):
    label counts agree with the mask, the un-opened run dominates it pixel-wise (an opening only removes), and sampled
    frames -- first, middle, last: every frame boundary case of the grid decode -- equal the oracle byte for byte."""
    n_frames, w, h, res, cap = 300, 1280, 720, 0.005, 1 << 21
    scan = Scan(ctx, pcs, n_frames, w, h)
    n_pts = n_frames * scan.ppf
    d_ann, d_mask = ctx.dev_alloc(n_pts * 32), ctx.dev_alloc(n_pts)
    vmap = ctx.create_map(res, cap)
    p_open = pcs.annotate_params(pcs.hsv_params(morph_open=1), lower=255, upper=255)
    ctx.annotate_frames_dev(scan.d_in, n_frames, w, h, scan.d_pose, p_open, vmap, d_ann, d_mask)
    keys, c0, c1, _ = vmap.export()
    mask = np.empty(n_pts, np.uint8)
    ctx.d2h(mask, d_mask)
    xs = np.empty(n_pts * 8, np.float32)
    ctx.d2h(xs, scan.d_in)
    fin = np.isfinite(xs[0::8])
    del xs
    assert set(np.unique(mask).tolist()) <= {0, 255}
    assert int(c1.sum(dtype=np.uint64)) == int(((mask == 255) & fin).sum())
    assert int(c0.sum(dtype=np.uint64)) + int(c1.sum(dtype=np.uint64)) == int(fin.sum())
    plain = np.empty(n_pts, np.uint8)
    ctx.annotate_frames_dev(scan.d_in, n_frames, w, h, scan.d_pose, pcs.annotate_params(lower=255, upper=255), None, None, d_mask)
    ctx.d2h(plain, d_mask)
    assert not ((mask == 255) & (plain == 0)).any() and int((plain == 255).sum()) > int((mask == 255).sum()) > 0
    op = O.default_params(morph_open=1)
    for f in (0, 1, n_frames // 2, n_frames - 1):
        o_ann, o_mask = O.annotate_frame(scan.frame(pcs, f), w, h, op)
        assert np.array_equal(mask[f * scan.ppf:(f + 1) * scan.ppf].reshape(h, w), o_mask), f
        assert scan.frame(pcs, f, d_ann).tobytes() == o_ann.tobytes(), f
    # map-only variant of the kernel: same map
    vmap.clear()
    ctx.annotate_frames_dev(scan.d_in, n_frames, w, h, scan.d_pose, p_open, vmap, None, None)
    assert all(np.array_equal(x, y) for x, y in zip(vmap.export()[:3], (keys, c0, c1)))
    vmap.destroy()
    ctx.dev_free(d_ann)
    ctx.dev_free(d_mask)
    scan.free()


def test_mesh_mask_2m_triangles_properties(ctx, pcs, O):
    """configs[3]-shaped: 2 M-triangle height-field against an aggregated 5 mm map; INSIDE + OUTSIDE partition the mesh,
    flags agree with the oracle on a sampled sub-mesh."""
    n_frames, w, h, res = 40, 1280, 720, 0.005
    scan = Scan(ctx, pcs, n_frames, w, h)
    vmap = ctx.create_map(res, 1 << 20)
    ctx.annotate_frames_dev(scan.d_in, n_frames, w, h, scan.d_pose, pcs.annotate_params(lower=255, upper=255), vmap, None, None)
    nx = ny = 1000
    xs = np.linspace(-0.45, 0.004 * n_frames + 0.45, nx + 1)
    ys = np.linspace(-0.25, 0.25, ny + 1)
    gx, gy = np.meshgrid(xs, ys, indexing="xy")
    verts = np.stack([gx.ravel(), gy.ravel(), 0.02 * np.sin(8 * gx.ravel()) * np.cos(6 * gy.ravel())], 1)
    idx = np.arange((nx + 1) * (ny + 1), dtype=np.int32).reshape(ny + 1, nx + 1)
    a, b, c, d = idx[:-1, :-1].ravel(), idx[:-1, 1:].ravel(), idx[1:, 1:].ravel(), idx[1:, :-1].ravel()
    tris = np.stack([np.stack([a, b, c], 1), np.stack([a, c, d], 1)], 1).reshape(-1, 3)
    faces = np.concatenate([np.full((tris.shape[0], 1), 3, np.int32), tris], 1).astype(np.int32)
    assert faces.shape[0] == 2_000_000
    iv, iff, _, flags = ctx.mesh_mask(vmap, verts, faces, pcs.RETURN_INSIDE)
    ov, of, _, flags2 = ctx.mesh_mask(vmap, verts, faces, pcs.RETURN_OUTSIDE)
    assert np.array_equal(flags, flags2)
    assert iff.shape[0] + of.shape[0] == faces.shape[0] and iff.shape[0] == int(flags.sum())
    assert 0.05 < flags.mean() < 0.99
    # emitted vertices are the un-welded corners of the kept triangles, in input order
    kept = np.nonzero(flags)[0]
    assert np.array_equal(iv.reshape(-1, 3, 3), verts[faces[kept, 1:]])
    assert np.array_equal(of[:, 1:], np.arange(3 * of.shape[0], dtype=np.int32).reshape(-1, 3))
    # oracle on every 97th triangle
    keys, c0, c1, _ = vmap.export()
    sample = np.arange(0, faces.shape[0], 97)
    o_flags = O.mesh_contacts_keys(keys[c1 > 0], res, verts, faces[sample])
    assert np.array_equal(flags[sample], o_flags)
    vmap.destroy()
    scan.free()
