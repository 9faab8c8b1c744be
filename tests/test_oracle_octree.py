"""Oracle octree: octomap update / prune rules, log-odds LUT, .bt codec pinned on the reference fixture."""
import hashlib
import os

import numpy as np
import pytest

REF_BT = "/root/reference/pcs_scan_integration/test/box_2m.bt"


def test_logodds_lut_bit_patterns(O):
    # float32 value after c hits
    want = [0x00000000, 0x3F58E883, 0x3FD8E883, 0x4022AE62, 0x4058E883, 0x4060B4BA, 0x4060B4BA, 0x4060B4BA]
    got = [int(np.float32(O.logodds_after_hits(c)).view(np.uint32)) for c in range(8)]
    assert got == want


def test_key_formula_and_range(O):
    import ctypes as C

    k = C.c_uint16(0)
    L = O.lib()
    assert L.orc_coord_to_key(100.0, 0.0, C.byref(k)) == 1 and k.value == 32768
    assert L.orc_coord_to_key(100.0, -0.005, C.byref(k)) == 1 and k.value == 32767
    assert L.orc_coord_to_key(100.0, 327.675, C.byref(k)) == 1 and k.value == 65535
    assert L.orc_coord_to_key(100.0, 327.68, C.byref(k)) == 0
    assert L.orc_coord_to_key(100.0, -327.68, C.byref(k)) == 1 and k.value == 0
    assert L.orc_coord_to_key(100.0, -327.69, C.byref(k)) == 0
    for bad in (float("nan"), float("inf"), float("-inf")):
        assert L.orc_coord_to_key(100.0, bad, C.byref(k)) == 0


def test_tree_equals_flat_counts(O):
    w, h, res = 160, 120, 0.005
    frames = [O.synth_frame(f, w, h) for f in range(3)]
    poses = [O.synth_pose(f) for f in range(3)]
    tree = O.Octree(res)
    ks, c1s = {}, {}
    for fr, po in zip(frames, poses):
        ann, _ = O.annotate_frame(fr, w, h)
        world = O.transform_points(po, ann)
        tree.insert_cloud(O.color_passthrough(world, 255, 255))
        k, c0, c1 = O.count_voxels(ann, res, po, 255, 255)
        for kk, a in zip(k.tolist(), c1.tolist()):
            c1s[kk] = c1s.get(kk, 0) + a
    occ = np.array(sorted(k for k, v in c1s.items() if v > 0), np.uint64)
    keys, vals = tree.leaf_voxels()
    assert np.array_equal(keys, occ)
    lut = np.array([O.logodds_after_hits(c) for c in range(6)], np.float32)
    exp = lut[np.minimum([c1s[int(k)] for k in occ], 5)]
    assert np.array_equal(vals.view(np.uint32), exp.view(np.uint32))
    n_before = tree.num_nodes()
    tree.finalize_tesseract()
    keys2, _ = tree.leaf_voxels()
    assert np.array_equal(keys2, occ)  # pruning never changes the occupied set
    assert tree.num_nodes() <= n_before


def test_octet_collapses_when_complete(O):
    t = O.Octree(0.1)
    for dx in range(2):
        for dy in range(2):
            for dz in range(2):
                t.insert_key(40000 + dx, 40000 + dy, 40000 + dz)
    kx, ky, kz, d, lo = t.leaves()
    # 8 equal-valued siblings auto-prune during insertion (octomap pruneNode on the way back up)
    assert len(kx) == 1 and d[0] == 15 and (kx[0], ky[0], kz[0]) == (40000, 40000, 40000)
    # a further hit expands the pruned node again and ends with 7 + 1 distinct values
    t.insert_key(40001, 40001, 40001)
    kx, ky, kz, d, lo = t.leaves()
    assert len(kx) == 8 and set(d.tolist()) == {16}
    assert sorted(set(lo.view(np.uint32).tolist())) == [0x3F58E883, 0x3FD8E883]
    t.finalize_tesseract()  # tesseract's prune: all children occupied leaves -> collapse, value of child 0
    kx, ky, kz, d, lo = t.leaves()
    assert len(kx) == 1 and d[0] == 15 and np.float32(lo[0]) == O.logodds_after_hits(1)


def test_bt_fixture_decodes_to_eight_unit_cubes(O, ref_fixtures):
    g = ref_fixtures
    assert float(g["bt_resolution"]) == 0.5 and int(g["bt_num_nodes"]) == 121
    assert len(g["bt_kx"]) == 8 and set(g["bt_depth"].tolist()) == {15}
    lo = (g["bt_kx"].astype(int) - 32768) * 0.5
    assert sorted(set(lo.tolist())) == [-1.0, 0.0]


def test_bt_writer_regenerates_reference_fixture_bytes(O, ref_fixtures):
    g = ref_fixtures
    t = O.Octree(float(g["bt_resolution"]))
    for x, y, z, d in zip(g["bt_kx"], g["bt_ky"], g["bt_kz"], g["bt_depth"]):
        s = 1 << (16 - int(d))
        for dx in range(s):
            for dy in range(s):
                for dz in range(s):
                    t.insert_key(int(x) + dx, int(y) + dy, int(z) + dz)
    data = t.write_bt()
    assert len(data) == int(g["bt_size"])
    assert hashlib.sha256(data).hexdigest() == str(g["bt_sha256"])
    again = O.Octree.read_bt(data)
    assert again.write_bt() == data


@pytest.mark.skipif(not os.path.exists(REF_BT), reason="reference checkout not present (GPU box)")
def test_bt_round_trip_on_reference_file(O):
    raw = open(REF_BT, "rb").read()
    assert O.Octree.read_bt(raw).write_bt() == raw
