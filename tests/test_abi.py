"""The C-ABI library loads and exports every symbol include/pcs_b200.h declares (no compute calls: runs on CPU)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def built():
    from point_cloud_segmentation_b200 import _build

    return _build.build()


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "pcs_b200.h")).read()
    return sorted(set(re.findall(r"PCS_API\s+[\w\s\*]+?\b(pcs_\w+)\s*\(", text)))


def test_header_symbols_all_exported(built):
    lib = ctypes.CDLL(built)
    names = declared_symbols()
    assert len(names) >= 40
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/pcs_b200.h but not exported"


def test_binding_table_matches_header(built):
    from point_cloud_segmentation_b200 import capi

    assert sorted(capi.SIGNATURES) == declared_symbols()
    capi.load_library()


def test_no_torch_or_oracle_in_product():
    """The product path must not import the oracle (test infrastructure) nor route through torch."""
    pkg = os.path.join(ROOT, "point_cloud_segmentation_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h", ".hpp")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle", src, re.M), f
                assert "pcs_oracle.h" not in src and "libpcs_oracle" not in src, f


def test_pod_layouts():
    from point_cloud_segmentation_b200 import capi

    assert capi.POINT_DTYPE.itemsize == 32 and capi.POINT_DTYPE.fields["b"][1] == 16  # sizeof(pcl::PointXYZRGB)
    assert ctypes.sizeof(capi.HsvParams) == 32 and ctypes.sizeof(capi.AnnotateParams) == 48


def test_version_and_defaults_without_gpu(built):
    from point_cloud_segmentation_b200 import capi

    assert capi.load_library().pcs_version() == 100
    p = capi.hsv_params()
    assert (p.h_lo, p.h_hi, p.s_lo, p.s_hi, p.v_lo, p.v_hi) == (95, 115, 95, 220, 40, 150)  # hsv_thresholding.h:47-52
    a = capi.annotate_params()
    assert (a.lower, a.upper, a.limit_negative) == (0, 255, 0)  # octomap_mesh_masking.h:107-109
    assert capi.mix64(0) == 0 and capi.mix64(1) != 1


def test_fails_loudly_without_cuda(built):
    """On a box without a GPU the product refuses to run instead of falling back to anything."""
    from point_cloud_segmentation_b200 import capi

    lib = capi.load_library()
    h = ctypes.c_void_p()
    rc = lib.pcs_ctx_create(0, None, ctypes.byref(h))
    if rc != 0:
        assert rc == capi.PCS_ERR_CUDA and b"no CPU fallback" in lib.pcs_last_error()
    else:  # running on the GPU box
        lib.pcs_ctx_destroy(h)


def test_cmake_project_configures_with_reference_target_names(tmp_path):
    """CMakeLists.txt exports pcs::pcs_detection_* / pcs::pcs_scan_integration_* like the reference
    (pcs_detection/CMakeLists.txt:14,35; pcs_scan_integration/CMakeLists.txt:14) -- configure only, no build."""
    import shutil
    import subprocess

    if not shutil.which("cmake") or not shutil.which("nvcc"):
        pytest.skip("cmake / nvcc not available")
    r = subprocess.run(["cmake", "-S", ROOT, "-B", str(tmp_path), "-DENABLE_TESTS=ON"], capture_output=True, text=True,
                       timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    text = open(os.path.join(ROOT, "CMakeLists.txt")).read()
    for tgt in ("pcs_detection_hsv_thresholding", "pcs_detection_point_cloud_annotator",
                "pcs_scan_integration_octomap_mesh_masking"):
        assert tgt in text
    assert "NAMESPACE pcs::" in text and "CMAKE_CUDA_ARCHITECTURES 100a" in text


def test_cmake_packages_are_found_by_a_reference_style_consumer(built, tmp_path):
    """The drop-in claim of INTEGRATION.md section 4: a consumer that does exactly what pcs_ros/CMakeLists.txt:6-7,48-64
    does -- find_package(pcs_scan_integration REQUIRED), find_package(pcs_detection REQUIRED), link the pcs:: targets --
    configures, compiles and links against an install prefix of this repository.  (libpcs_b200.so itself is the one the
    Python-driven build made: cmake --install only copies files, so the CUDA sources are not compiled a second time.)"""
    import shutil
    import subprocess

    if not shutil.which("cmake") or not shutil.which("nvcc"):
        pytest.skip("cmake / nvcc not available")
    from point_cloud_segmentation_b200 import _build

    build, prefix, consumer = tmp_path / "build", tmp_path / "prefix", tmp_path / "consumer"
    r = subprocess.run(["cmake", "-S", ROOT, "-B", str(build)], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    shutil.copy(_build.LIB_PATH, build / "libpcs_b200.so")
    r = subprocess.run(["cmake", "--install", str(build), "--prefix", str(prefix)], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    for pkg in ("pcs_b200", "pcs_detection", "pcs_scan_integration"):
        for f in (f"{pkg}-config.cmake", f"{pkg}-config-version.cmake", f"{pkg}-targets.cmake"):
            assert (prefix / "lib" / "cmake" / pkg / f).exists(), (pkg, f)
    consumer.mkdir()
    (consumer / "CMakeLists.txt").write_text("""cmake_minimum_required(VERSION 3.5.0)
project(pcs_ros_like CXX)
find_package(pcs_scan_integration REQUIRED)
find_package(pcs_detection REQUIRED)
add_executable(node node.cpp)
target_link_libraries(node pcs::pcs_detection_point_cloud_annotator pcs::pcs_detection_hsv_thresholding
                      pcs::pcs_scan_integration_octomap_mesh_masking)
set_target_properties(node PROPERTIES CXX_STANDARD 17)
""")
    (consumer / "node.cpp").write_text("""#include <pcs_detection/hsv_thresholding.h>
#include <pcs_detection/point_cloud_annotator.h>
#include <pcs_detection/utils.h>
#include <pcs_scan_integration/octomap_mesh_masking.h>
int main()
{
  pcs_scan_integration::OctomapMeshMask masker;
  pcs_compat::Mat a(1, 1, pcs_compat::CV_8UC1), b(1, 1, pcs_compat::CV_8UC1), c;
  a.data[0] = 3;
  b.data[0] = 5;
  return pcs_detection::applyMask(a, b, c) && c.data[0] == 15 && pcs_version() == PCS_VERSION ? 0 : 1;
}
""")
    cb = tmp_path / "consumer_build"
    r = subprocess.run(["cmake", "-S", str(consumer), "-B", str(cb), f"-DCMAKE_PREFIX_PATH={prefix}"], capture_output=True,
                       text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    r = subprocess.run(["cmake", "--build", str(cb)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    env = dict(os.environ, LD_LIBRARY_PATH=str(prefix / "lib") + ":" + os.environ.get("LD_LIBRARY_PATH", ""))
    r = subprocess.run([str(cb / "node")], capture_output=True, text=True, timeout=60, env=env)
    assert r.returncode == 0, r.stdout[-1000:] + r.stderr[-1000:]  # applyMask and pcs_version need no GPU
