"""C++ drop-in adapters (point_cloud_segmentation_b200/cpp/include): they compile on CPU, and on a GPU the reference's
gtests re-expressed against them pass (cpp/test/*.cpp mirror pcs_detection/test/*.cpp and
pcs_scan_integration/test/octomap_mesh_masking_unit.cpp)."""
import os
import subprocess

import numpy as np
import pytest

from util import faces4_of


@pytest.fixture(scope="module")
def binaries():
    from point_cloud_segmentation_b200 import _build
    from point_cloud_segmentation_b200.cpp import build_cpp

    _build.build()
    return build_cpp.build()


def test_adapters_compile(binaries):
    assert all(os.path.exists(b) for b in binaries)


def write_fixtures(O, g, d):
    """plane_4m.stl / box_2m.ply / box_2m.bt regenerated from tests/golden (the .bt byte-identical to the reference's)."""
    v, f = g["plane_verts"], g["plane_faces"]
    rec = np.zeros(f.shape[0], np.dtype([("n", "<f4", 3), ("v", "<f4", (3, 3)), ("attr", "<u2")]))
    rec["v"] = v[f]
    rec["n"] = (0, 0, 1)
    with open(os.path.join(d, "plane_4m.stl"), "wb") as fh:
        fh.write(b"regenerated from tests/golden".ljust(80, b" ") + np.uint32(f.shape[0]).tobytes() + rec.tobytes())
    with open(os.path.join(d, "box_2m.ply"), "wb") as fh:
        fh.write(O.write_ply_ascii(g["box_verts"].astype(np.float64), None, faces4_of(g["box_faces"])))
    t = O.Octree(float(g["bt_resolution"]))
    for x, y, z, dep in zip(g["bt_kx"], g["bt_ky"], g["bt_kz"], g["bt_depth"]):
        s = 1 << (16 - int(dep))
        for dx in range(s):
            for dy in range(s):
                for dz in range(s):
                    t.insert_key(int(x) + dx, int(y) + dy, int(z) + dz)
    with open(os.path.join(d, "box_2m.bt"), "wb") as fh:
        fh.write(t.write_bt())


@pytest.mark.gpu
def test_detection_unit(binaries):
    r = subprocess.run([binaries[0]], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "0 failures" in r.stdout


@pytest.mark.gpu
def test_octomap_mesh_masking_unit(binaries, O, ref_fixtures, tmp_path):
    write_fixtures(O, ref_fixtures, str(tmp_path))
    env = dict(os.environ, PCS_TEST_DATA_DIR=str(tmp_path))
    r = subprocess.run([binaries[1]], capture_output=True, text=True, timeout=300, env=env)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "0 failures" in r.stdout
    # saveMaskedMesh wrote ASCII PLYs the oracle's loader can read back
    v, f = O.load_ply_ascii(open(os.path.join(tmp_path, "test_output_RETURN_OUTSIDE.ply"), "rb").read())
    assert f.shape[0] == 11400
