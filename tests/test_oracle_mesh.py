"""Oracle mesh mask pinned on the reference goldens (box_2m.bt x plane_4m.stl) and mesh I/O."""
import os

import numpy as np
import pytest

from util import faces4_of, grid_mesh

REF = "/root/reference/pcs_scan_integration/test"


def fixture_tree(O, g):
    t = O.Octree(float(g["bt_resolution"]))
    for x, y, z, d in zip(g["bt_kx"], g["bt_ky"], g["bt_kz"], g["bt_depth"]):
        s = 1 << (16 - int(d))
        for dx in range(s):
            for dy in range(s):
                for dz in range(s):
                    t.insert_key(int(x) + dx, int(y) + dy, int(z) + dz)
    t.finalize_tesseract()
    return t


def test_maskmesh_goldens(O, ref_fixtures):
    g = ref_fixtures
    verts = g["plane_verts"].astype(np.float64)
    faces = faces4_of(g["plane_faces"])
    assert verts.shape[0] == int(g["golden_plane_nv"]) == 6273 and faces.shape[0] == int(g["golden_plane_nt"]) == 12288
    tree = fixture_tree(O, g)
    kx, ky, kz, d, lo = tree.leaves()
    assert len(kx) == 8 and set(d.tolist()) == {15}  # pruned back to the 8 unit cubes of the fixture
    pairs = tree.mesh_contacts(verts, faces)
    flags = (pairs > 0).astype(np.uint8)
    # octomap_mesh_masking_unit.cpp:182-183: RETURN_OUTSIDE -> 11400 triangles / 34200 vertices
    ov, of, rgb = O.mask_mesh_emit(verts, faces, flags, 1)
    assert of.shape[0] == int(g["golden_outside_tris"]) == 11400 and ov.shape[0] == 34200
    # :158-159 RETURN_COLORIZED -> the whole mesh
    cv, cf, crgb = O.mask_mesh_emit(verts, faces, flags, 2)
    assert cv.shape[0] == 6273 and cf.shape[0] == 12288
    # :170-171 RETURN_INSIDE is Bullet's contact multiset (1975); geometry: 888 unique triangles, 1976 pairs
    iv, iff, _ = O.mask_mesh_emit(verts, faces, flags, 0)
    assert iff.shape[0] == 888 and iv.shape[0] == 2664
    assert int(pairs.sum()) == 1976 and abs(int(pairs.sum()) - int(g["golden_inside_entries_bullet"])) == 1
    # flat-key path (finest voxels) gives the same triangle set: the union of children equals the pruned leaf
    keys, _ = tree.leaf_voxels()
    assert keys.size == 64
    assert np.array_equal(O.mesh_contacts_keys(keys, tree.resolution, verts, faces), flags)


def test_emit_layout(O):
    verts, faces = grid_mesh(2, 1, 0, 2, 0, 1)
    flags = np.array([1, 0, 0, 1], np.uint8)
    ov, of, rgb = O.mask_mesh_emit(verts, faces, flags, 0)
    assert of.tolist() == [[3, 0, 1, 2], [3, 3, 4, 5]]
    assert np.array_equal(ov[:3], verts[faces[0, 1:]]) and np.array_equal(ov[3:], verts[faces[3, 1:]])
    assert (rgb == [0, 128, 0]).all()
    ov, of, rgb = O.mask_mesh_emit(verts, faces, flags, 2)
    red = np.zeros(verts.shape[0], bool)
    red[faces[0, 1:]] = True
    red[faces[3, 1:]] = True
    assert (rgb[red] == [255, 0, 0]).all() and (rgb[~red] == [0, 128, 0]).all()


def test_sat_touching_counts_and_separation(O):
    lo, hi = [0, 0, 0], [1, 1, 1]
    tri = lambda z: ([0.2, 0.2, z], [0.8, 0.2, z], [0.2, 0.8, z])
    assert O.tri_box_overlap(lo, hi, *tri(0.5))
    assert O.tri_box_overlap(lo, hi, *tri(1.0))       # lies in the top face: distance 0 is a contact
    assert not O.tri_box_overlap(lo, hi, *tri(1.0 + 1e-12))
    assert O.tri_box_overlap(lo, hi, [1, 1, 1], [2, 1, 1], [1, 2, 1])  # touches one corner
    # diagonal triangle whose AABB overlaps the box but which misses it (needs the cross / normal axes)
    assert not O.tri_box_overlap(lo, hi, [1.6, -0.4, 0.5], [2.4, 0.4, 0.5], [1.6, 0.4, 2.0])
    assert not O.tri_box_overlap(lo, hi, [3.5, -1, 0.5], [-1, 3.5, 0.5], [3, 3, 0.5])  # edge x + y = 2.5 misses (1,1)
    assert O.tri_box_overlap(lo, hi, [3.01, -1, 0.5], [-1, 3.01, 0.5], [3, 3, 0.5]) is False
    assert O.tri_box_overlap(lo, hi, [3.0, -1, 0.5], [-1, 3.0, 0.5], [3, 3, 0.5]) is True  # edge x + y = 2 grazes the corner (1,1)


def test_keys_path_matches_brute_force(O):
    rng = np.random.default_rng(5)
    res = 0.05
    kk = rng.integers(32768 - 6, 32768 + 6, (300, 3))
    kk = np.unique(kk, axis=0)
    keys = (kk[:, 2].astype(np.uint64) << np.uint64(32)) | (kk[:, 1].astype(np.uint64) << np.uint64(16)) | kk[:, 0].astype(np.uint64)
    lo = (kk - 32768) * res
    hi = (kk - 32767) * res
    verts = rng.uniform(-0.4, 0.4, (600, 3))
    # some vertices exactly on voxel faces
    verts[::7] = np.round(verts[::7] / res) * res
    tris = rng.integers(0, 600, (400, 3))
    near = rng.integers(0, 600, 400)
    verts2 = np.concatenate([verts, verts[near] + rng.uniform(-0.06, 0.06, (400, 3))])
    tris[:, 1] = near
    tris[:, 2] = 600 + np.arange(400)
    faces = faces4_of(tris)
    brute = O.mesh_contacts_brute(lo, hi, verts2, faces)
    flags = O.mesh_contacts_keys(keys, res, verts2, faces)
    assert np.array_equal(flags, (brute > 0).astype(np.uint8))
    assert 0 < flags.sum() < flags.size


def test_stl_ply_writers_and_loaders(O, ref_fixtures, tmp_path):
    g = ref_fixtures
    # re-create a binary STL from the welded golden mesh, load + weld it again
    v = g["plane_verts"]
    f = g["plane_faces"]
    rec = np.zeros(f.shape[0], np.dtype([("n", "<f4", 3), ("v", "<f4", (3, 3)), ("attr", "<u2")]))
    rec["v"] = v[f]
    data = b"\0" * 80 + np.uint32(f.shape[0]).tobytes() + rec.tobytes()
    lv, lf = O.load_stl_binary(data)
    assert lv.shape == (6273, 3) and lf.shape == (12288, 4)
    assert np.array_equal(lv[lf[:, 1:]].astype(np.float32), v[f])
    text = O.write_ply_ascii(g["box_verts"].astype(np.float64), None, faces4_of(g["box_faces"]))
    pv, pf = O.load_ply_ascii(text)
    assert pv.shape == (8, 3) and pf.shape == (12, 4)
    colored = O.write_ply_ascii(pv, np.tile(np.array([0, 128, 0], np.uint8), (8, 1)), pf).decode()
    assert "property uchar red" in colored and colored.splitlines()[-1].startswith("3 ")


@pytest.mark.skipif(not os.path.exists(REF), reason="reference checkout not present (GPU box)")
def test_loaders_on_reference_files(O):
    v, f = O.load_stl_binary(open(os.path.join(REF, "plane_4m.stl"), "rb").read())
    assert v.shape[0] == 6273 and f.size == 12288 * 4  # octomap_mesh_masking_unit.cpp:76-77
    v, f = O.load_ply_ascii(open(os.path.join(REF, "box_2m.ply"), "rb").read())
    assert v.shape[0] == 8 and f.size == 12 * 4  # :85-86


def test_far_vertices_terminate(O):
    """Vertices at 1e6 .. DBL_MAX metres must not spin the key-range walk (bounded by clamping to the key range)."""
    from test_gpu_mesh import far_vertex_case

    res, keys, verts, faces = far_vertex_case()
    flags = O.mesh_contacts_keys(keys, res, verts, faces)
    assert flags.tolist() == [0, 0, 0, 1, 0, 1, 1, 0]
