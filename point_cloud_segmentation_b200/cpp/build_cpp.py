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


if __name__ == "__main__":
    print("\n".join(build(force=True)))
