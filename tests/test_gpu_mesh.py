"""Stage 4 parity through the C ABI: maskMesh vs the reference goldens and the oracle.  Triangle set bit-exact."""
import numpy as np
import pytest

from util import faces4_of, grid_mesh, surface

pytestmark = pytest.mark.gpu


def fixture_map(ctx, g):
    """box_2m.bt as a voxel map: its 8 depth-15 leaves expanded to 64 voxels at the tree resolution (0.5 m)."""
    keys = []
    for x, y, z, d in zip(g["bt_kx"], g["bt_ky"], g["bt_kz"], g["bt_depth"]):
        s = 1 << (16 - int(d))
        for dx in range(s):
            for dy in range(s):
                for dz in range(s):
                    keys.append(((int(z) + dz) << 32) | ((int(y) + dy) << 16) | (int(x) + dx))
    keys = np.array(keys, np.uint64)
    vmap = ctx.create_map(float(g["bt_resolution"]), 1024)
    vmap.merge(keys, None, np.ones(keys.size, np.uint32))
    return vmap, keys


def test_maskmesh_reference_goldens(ctx, pcs, O, ref_fixtures):
    # pcs_scan_integration/test/octomap_mesh_masking_unit.cpp:137-185
    g = ref_fixtures
    verts = g["plane_verts"].astype(np.float64)
    faces = faces4_of(g["plane_faces"])
    vmap, keys = fixture_map(ctx, g)
    ov, of, rgb, flags = ctx.mesh_mask(vmap, verts, faces, pcs.RETURN_COLORIZED)
    assert ov.shape[0] == 6273 and of.size == 12288 * 4                      # :158-159
    ov, of, rgb, flags = ctx.mesh_mask(vmap, verts, faces, pcs.RETURN_OUTSIDE)
    assert ov.shape[0] == 34200 and of.size == 11400 * 4                     # :182-183
    ov, of, rgb, flags = ctx.mesh_mask(vmap, verts, faces, pcs.RETURN_INSIDE)
    assert of.shape[0] == 888 and ov.shape[0] == 2664                        # set semantics of :170-171 (Bullet: 1975 contacts)
    o_flags = O.mesh_contacts_keys(keys, float(g["bt_resolution"]), verts, faces)
    assert np.array_equal(flags, o_flags)
    for mt in (0, 1, 2):
        got = ctx.mesh_mask(vmap, verts, faces, mt)
        want = O.mask_mesh_emit(verts, faces, o_flags, mt)
        assert got[0].tobytes() == want[0].tobytes() and np.array_equal(got[1], want[1]) and np.array_equal(got[2], want[2])
    vmap.destroy()


@pytest.mark.parametrize("res,nx", [(0.01, 100), (0.005, 220)])
def test_heightfield_mesh_vs_oracle(ctx, pcs, O, res, nx):
    """Config-1-shaped: a 640x480-ish scan integrated at `res`, a 2*nx*nx/2-triangle height-field masked by it."""
    w, h = 320, 240
    frames = np.concatenate([O.synth_frame(f, w, h) for f in range(3)])
    poses = np.stack([O.synth_pose(f) for f in range(3)])
    vmap = ctx.create_map(res, 1 << 18)
    ctx.annotate_frames(frames, 3, w, h, poses, pcs.annotate_params(lower=255, upper=255), vmap, want_annotated=False,
                        want_mask=False)
    keys, c0, c1, _ = vmap.export()
    verts, faces = grid_mesh(nx, nx // 2, -0.45, 0.45, -0.25, 0.25, surface)
    o_flags = O.mesh_contacts_keys(keys[c1 > 0], res, verts, faces)
    assert 0.02 < o_flags.mean() < 0.9
    for mt in (pcs.RETURN_INSIDE, pcs.RETURN_OUTSIDE, pcs.RETURN_COLORIZED):
        ov, of, rgb, flags = ctx.mesh_mask(vmap, verts, faces, mt)
        assert np.array_equal(flags, o_flags)
        wv, wf, wrgb = O.mask_mesh_emit(verts, faces, o_flags, mt)
        assert ov.tobytes() == wv.tobytes() and np.array_equal(of, wf) and np.array_equal(rgb, wrgb)
    vmap.destroy()


def test_count0_only_voxels_do_not_mask(ctx, pcs, O):
    """Voxels that only hold points rejected by colorPassthrough are not part of the reference octree."""
    vmap = ctx.create_map(0.1, 64)
    key = np.array([(32768 << 32) | (32768 << 16) | 32768], np.uint64)
    vmap.merge(key, np.array([5], np.uint32), np.array([0], np.uint32))
    verts, faces = grid_mesh(2, 2, 0.0, 0.1, 0.0, 0.1, lambda x, y: 0.05 + 0 * x)
    _, of, _, flags = ctx.mesh_mask(vmap, verts, faces, pcs.RETURN_INSIDE)
    assert flags.sum() == 0 and of.shape[0] == 0
    vmap.merge(key, None, np.array([1], np.uint32))
    _, of, _, flags = ctx.mesh_mask(vmap, verts, faces, pcs.RETURN_INSIDE)
    assert flags.all() and of.shape[0] == 8
    vmap.destroy()


def test_large_triangles_take_the_voxel_major_path(ctx, pcs, O, ref_fixtures):
    """box_2m.ply (12 triangles, 2 m edges) against 2 cm voxels: key ranges of ~10^6 voxels per triangle."""
    g = ref_fixtures
    verts = g["box_verts"].astype(np.float64)
    faces = faces4_of(g["box_faces"])
    rng = np.random.default_rng(12)
    kk = rng.integers(32768 - 60, 32768 + 60, (4000, 3))
    kk[:500, 2] = 32768 + 49  # a layer just under the top face z = 1 (voxel [0.98, 1.00])
    kk[500:800, 0] = 32768 - 51  # just outside x = -1 (voxel [-1.02, -1.00]) -> touches the face
    keys = np.unique((kk[:, 2].astype(np.uint64) << np.uint64(32)) | (kk[:, 1].astype(np.uint64) << np.uint64(16)) |
                     kk[:, 0].astype(np.uint64))
    vmap = ctx.create_map(0.02, 1 << 14)
    vmap.merge(keys, None, np.ones(keys.size, np.uint32))
    o_flags = O.mesh_contacts_keys(keys, 0.02, verts, faces)
    _, _, _, flags = ctx.mesh_mask(vmap, verts, faces, pcs.RETURN_COLORIZED)
    assert np.array_equal(flags, o_flags) and 0 < flags.sum()
    vmap.destroy()


def test_vertices_on_voxel_faces_and_degenerate_triangles(ctx, pcs, O):
    rng = np.random.default_rng(13)
    res = 0.05
    kk = np.unique(rng.integers(32768 - 6, 32768 + 6, (300, 3)), axis=0)
    keys = (kk[:, 2].astype(np.uint64) << np.uint64(32)) | (kk[:, 1].astype(np.uint64) << np.uint64(16)) | kk[:, 0].astype(np.uint64)
    verts = rng.uniform(-0.4, 0.4, (900, 3))
    verts[::3] = np.round(verts[::3] / res) * res  # exactly on voxel faces / edges / corners
    tris = rng.integers(0, 900, (3000, 3))
    tris[::50, 2] = tris[::50, 1]  # degenerate (zero-area) triangles
    tris[::75] = tris[::75, :1]    # all three vertices identical
    # keep triangles small so that the walk stays local
    verts = np.concatenate([verts, verts[tris[:, 0]] + rng.uniform(-0.07, 0.07, (3000, 3)),
                            verts[tris[:, 0]] + rng.uniform(-0.07, 0.07, (3000, 3))])
    small = np.stack([tris[:, 0], 900 + np.arange(3000), 3900 + np.arange(3000)], 1)
    small[::50, 2] = small[::50, 1]
    small[::75] = small[::75, :1]
    faces = faces4_of(small)
    vmap = ctx.create_map(res, 4096)
    vmap.merge(keys, None, np.ones(keys.size, np.uint32))
    o_flags = O.mesh_contacts_keys(keys, res, verts, faces)
    _, _, _, flags = ctx.mesh_mask(vmap, verts, faces, pcs.RETURN_COLORIZED)
    assert np.array_equal(flags, o_flags)
    vmap.destroy()


def far_vertex_case():
    """Finite vertices far outside the key range: 1e6 m, 1e30 m, DBL_MAX; some triangles reach back in."""
    res = 0.005
    kk = np.array([[32768 + i, 32768, 32768] for i in range(-3, 4)] + [[65535, 32768, 32768], [0, 32768, 32768]])
    keys = (kk[:, 2].astype(np.uint64) << np.uint64(32)) | (kk[:, 1].astype(np.uint64) << np.uint64(16)) | kk[:, 0].astype(np.uint64)
    big = np.finfo(np.float64).max
    verts = np.array([
        [1e6, 1e6, 1e6], [1e6 + 1, 1e6, 1e6], [1e6, 1e6 + 1, 1e6],            # 0-2   far, all axes
        [1e30, 1e30, 1e30], [-1e30, 2e30, 1e30], [1e30, -1e30, 3e30],        # 3-5   very far, mixed signs
        [big, big, big], [-big, -big, big], [big, -big, big],                # 6-8   DBL_MAX (z stays outside the key range)
        [0.001, 0.001, 0.001], [1e30, 0.002, 0.002], [-1e30, 0.003, 0.001],  # 9-11  thin sliver along x through the map
        [0.001, 1e6, 0.001], [0.002, 2e6, 0.001], [0.001, 1.5e6, 0.002],     # 12-14 far on y only
        [0.0, 0.0, 0.0], [0.004, 0.0, 0.0], [0.0, 0.004, 0.004],             # 15-17 an ordinary hit
    ])
    tris = np.array([[0, 1, 2], [3, 4, 5], [6, 7, 8], [9, 10, 11], [12, 13, 14], [15, 16, 17], [9, 10, 10], [3, 3, 3]])
    return res, keys, verts, faces4_of(tris)


def test_far_vertices_terminate_and_match_the_oracle(ctx, pcs, O):
    res, keys, verts, faces = far_vertex_case()
    vmap = ctx.create_map(res, 1024)
    vmap.merge(keys, None, np.ones(keys.size, np.uint32))
    o_flags = O.mesh_contacts_keys(keys, res, verts, faces)
    _, _, _, flags = ctx.mesh_mask(vmap, verts, faces, pcs.RETURN_COLORIZED)
    assert np.array_equal(flags, o_flags)
    assert flags.tolist() == [0, 0, 0, 1, 0, 1, 1, 0]
    vmap.destroy()


def test_device_resident_mask_equals_host_call(ctx, pcs, O):
    """pcs_mesh_mask_dev (mesh and outputs in HBM) against pcs_mesh_mask and the oracle, all three mask types; also vertices
    exactly on voxel faces, where the per-vertex records widen the key box."""
    res = 0.01
    w, h = 320, 240
    frames = np.concatenate([O.synth_frame(f, w, h) for f in range(2)])
    poses = np.stack([O.synth_pose(f) for f in range(2)])
    vmap = ctx.create_map(res, 1 << 18)
    ctx.annotate_frames(frames, 2, w, h, poses, pcs.annotate_params(lower=255, upper=255), vmap, want_annotated=False,
                        want_mask=False)
    keys, c0, c1, _ = vmap.export()
    # y spacing 0.005 = half a voxel: every second row of vertices lies exactly on a voxel face
    verts, faces = grid_mesh(90, 200, -0.45, 0.45, -0.5, 0.5, surface)
    verts[:, 1] = np.round(verts[:, 1] / 0.005) * 0.005
    verts[::7, 2] = np.round(verts[::7, 2] / res) * res
    nv, nt = verts.shape[0], faces.shape[0]
    o_flags = O.mesh_contacts_keys(keys[c1 > 0], res, verts, faces)
    assert 0.05 < o_flags.mean() < 0.95
    d_v, d_f, d_fl = ctx.dev_alloc(verts.nbytes), ctx.dev_alloc(faces.nbytes), ctx.dev_alloc(nt)
    d_ov, d_of, d_rgb, d_n = ctx.dev_alloc(nt * 72), ctx.dev_alloc(nt * 16), ctx.dev_alloc(max(nt * 9, nv * 3)), ctx.dev_alloc(8)
    ctx.h2d(d_v, verts)
    ctx.h2d(d_f, faces)
    for mt in (pcs.RETURN_INSIDE, pcs.RETURN_OUTSIDE, pcs.RETURN_COLORIZED):
        ctx.mesh_mask_dev(vmap, d_v, nv, d_f, nt, mt, nt, d_ov, d_of, d_rgb, d_fl, d_n)
        flags, n_out = np.empty(nt, np.uint8), np.zeros(1, np.uint64)
        ctx.d2h(flags, d_fl)
        ctx.d2h(n_out, d_n)
        assert np.array_equal(flags, o_flags)
        wv, wf, wrgb = O.mask_mesh_emit(verts, faces, o_flags, mt)
        hv, hf, hrgb, hflags = ctx.mesh_mask(vmap, verts, faces, mt)
        assert np.array_equal(hflags, o_flags) and hv.tobytes() == wv.tobytes() and np.array_equal(hf, wf) and np.array_equal(hrgb, wrgb)
        assert int(n_out[0]) == wf.shape[0]
        if mt == pcs.RETURN_COLORIZED:
            rgb = np.empty((nv, 3), np.uint8)
            ctx.d2h(rgb, d_rgb)
            assert np.array_equal(rgb, wrgb)
        else:
            k = wf.shape[0]
            ov, of, rgb = np.empty((3 * k, 3), np.float64), np.empty((k, 4), np.int32), np.empty((3 * k, 3), np.uint8)
            if k:
                ctx.d2h(ov, d_ov)
                ctx.d2h(of, d_of)
                ctx.d2h(rgb, d_rgb)
            assert ov.tobytes() == wv.tobytes() and np.array_equal(of, wf) and np.array_equal(rgb, wrgb)
    # count-only form
    ctx.mesh_mask_dev(vmap, d_v, nv, d_f, nt, pcs.RETURN_INSIDE, 0, None, None, None, d_fl, d_n)
    n_out = np.zeros(1, np.uint64)
    ctx.d2h(n_out, d_n)
    assert int(n_out[0]) == int(o_flags.sum())
    for p in (d_v, d_f, d_fl, d_ov, d_of, d_rgb, d_n):
        ctx.dev_free(p)
    vmap.destroy()


def test_empty_mesh_and_bad_mask_type(ctx, pcs):
    vmap = ctx.create_map(0.1, 64)
    ov, of, rgb, flags = ctx.mesh_mask(vmap, np.zeros((0, 3)), np.zeros((0, 4), np.int32), pcs.RETURN_OUTSIDE)
    assert ov.shape[0] == 0 and of.shape[0] == 0
    verts, faces = grid_mesh(1, 1, 0, 1, 0, 1)
    with pytest.raises(pcs.PcsError) as e:  # octomap_mesh_mask_node.cpp:93-98 "Invalid mask type"
        ctx.mesh_mask(vmap, verts, faces, 3)
    assert "Invalid mask type" in str(e.value)
    ov, of, rgb, flags = ctx.mesh_mask(vmap, verts, faces, pcs.RETURN_OUTSIDE)  # empty map: everything is outside
    assert of.shape[0] == 2 and flags.sum() == 0
    vmap.destroy()


def test_python_mirror_of_octomap_mesh_mask(ctx, pcs, O, ref_fixtures, tmp_path):
    """scan_integration.OctomapMeshMask follows the reference's unit test (octomap_mesh_masking_unit.cpp:45-185)."""
    from point_cloud_segmentation_b200.scan_integration import OctomapMeshMask, colorPassthrough
    from util import ramp_cloud

    g = ref_fixtures
    masker = OctomapMeshMask(ctx)
    assert masker.getOctree() is None and masker.getInputMesh() is None and masker.getMaskedMesh() is None
    assert not masker.maskMesh(OctomapMeshMask.MaskType.RETURN_INSIDE)
    cloud = ramp_cloud(pcs.POINT_DTYPE)
    masker.setOctree(cloud, 0.1)
    assert masker.getOctree() is not None
    assert colorPassthrough(cloud, 25, 100, False, ctx).size == 76
    vmap, _ = fixture_map(ctx, g)
    masker.setOctree(vmap)
    masker.setInputMesh(g["plane_verts"].astype(np.float64), faces4_of(g["plane_faces"]))
    assert masker.getInputMesh()[0].shape[0] == 6273
    assert masker.maskMesh(OctomapMeshMask.MaskType.RETURN_COLORIZED) and masker.getMaskedMesh()[1].shape[0] == 12288
    assert masker.maskMesh(OctomapMeshMask.MaskType.RETURN_OUTSIDE) and masker.getMaskedMesh()[1].shape[0] == 11400
    path = str(tmp_path / "masked_mesh.ply")
    assert masker.saveMaskedMesh(path)
    v, f = O.load_ply_ascii(open(path, "rb").read())
    assert f.shape[0] == 11400
    assert masker.maskMesh(1) and not masker.maskMesh(7)  # the action's integers; invalid type -> False
    vmap.destroy()
