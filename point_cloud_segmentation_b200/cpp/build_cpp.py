"""Builds the C++ adapter tests (g++, C++17) against libpcs_b200.so into cpp/build/."""
from __future__ import annotations

import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
_PKG = os.path.dirname(_HERE)
_ROOT = os.path.dirname(_PKG)
BUILD = os.path.join(_HERE, "build")
TESTS = ["detection_unit", "octomap_mesh_masking_unit"]


def build(force: bool = False) -> list[str]:
    os.makedirs(BUILD, exist_ok=True)
    lib = os.path.join(_PKG, "libpcs_b200.so")
    if not os.path.exists(lib):
        raise RuntimeError("build libpcs_b200.so first (point_cloud_segmentation_b200._build.build())")
    headers = []
    for d, _, files in os.walk(os.path.join(_HERE, "include")):
        headers += [os.path.join(d, f) for f in files]
    headers += [os.path.join(_ROOT, "include", "pcs_b200.h"), os.path.join(_HERE, "test", "mini_test.h")]
    newest = max(os.path.getmtime(h) for h in headers)
    outs = []
    for t in TESTS:
        src = os.path.join(_HERE, "test", t + ".cpp")
        out = os.path.join(BUILD, t)
        outs.append(out)
        if not force and os.path.exists(out) and os.path.getmtime(out) >= max(newest, os.path.getmtime(src)):
            continue
        cmd = ["g++", "-std=c++17", "-O2", "-Wall", "-Wextra", "-Wno-unused-parameter", "-I", os.path.join(_ROOT, "include"),
               "-I", os.path.join(_HERE, "include"), "-I", os.path.join(_HERE, "test"), src, "-o", out, lib,
               "-Wl,-rpath," + _PKG, "-lpthread"]
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError(f"g++ failed for {t}:\n{res.stdout}\n{res.stderr}")
    return outs


def build_pybind(force: bool = False) -> str:
    """The pybind11 module for the Python detection scripts: cpp/pybind/pcs_detection_py.cpp -> pcs_detection_py.<abi>.so
    next to libpcs_b200.so (found through an $ORIGIN rpath).  Needs the Python headers and pybind11."""
    import sysconfig

    import pybind11

    lib = os.path.join(_PKG, "libpcs_b200.so")
    src = os.path.join(_HERE, "pybind", "pcs_detection_py.cpp")
    out = os.path.join(_PKG, "pcs_detection_py" + sysconfig.get_config_var("EXT_SUFFIX"))
    hdr = os.path.join(_ROOT, "include", "pcs_b200.h")
    if not force and os.path.exists(out) and os.path.getmtime(out) >= max(os.path.getmtime(src), os.path.getmtime(hdr)):
        return out
    cmd = ["g++", "-std=c++17", "-O2", "-shared", "-fPIC", "-fvisibility=hidden", "-I", sysconfig.get_paths()["include"],
           "-I", pybind11.get_include(), "-I", os.path.join(_ROOT, "include"), src, "-o", out, lib, "-Wl,-rpath,$ORIGIN"]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError(f"g++ failed for the pybind11 module:\n{res.stdout}\n{res.stderr}")
    return out


if __name__ == "__main__":
    print("\n".join(build(force=True)))
    print(build_pybind(force=True))
