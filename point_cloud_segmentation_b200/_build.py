"""In-tree nvcc build of libpcs_b200.so (sm_100a only).

The library is built next to the sources so that it travels with a repository snapshot; nothing is installed into
site-packages and nothing is JIT-cached under ~/.cache.  nvcc cross-compiles without a GPU.
"""
from __future__ import annotations

import os
import shutil
import subprocess
from concurrent.futures import ThreadPoolExecutor

_PKG = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_PKG)
CSRC = os.path.join(_PKG, "csrc")
INCLUDE = os.path.join(_ROOT, "include")
BUILD_DIR = os.path.join(CSRC, "build")
LIB_PATH = os.path.join(_PKG, "libpcs_b200.so")

ARCH_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON_FLAGS = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden",
                "-I", INCLUDE, "-I", CSRC]
# files whose floating-point results are compared bit-for-bit with the oracle: no FMA contraction
NO_FMAD = {"mesh_mask.cu", "synth.cu"}


def nvcc() -> str:
    exe = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(exe):
        raise RuntimeError("nvcc not found: libpcs_b200 needs the CUDA toolkit to build")
    return exe


def sources() -> list[str]:
    return sorted(f for f in os.listdir(CSRC) if f.endswith(".cu"))


def _deps_mtime() -> float:
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    deps += [os.path.join(INCLUDE, f) for f in os.listdir(INCLUDE)]
    deps.append(os.path.abspath(__file__))
    return max(os.path.getmtime(d) for d in deps)


def _compile(src: str, verbose: bool, extra: list[str]) -> str:
    obj = os.path.join(BUILD_DIR, src[:-3] + ".o")
    cmd = [nvcc(), *ARCH_FLAGS, *COMMON_FLAGS, *extra]
    if src in NO_FMAD:
        cmd.append("-fmad=false")
    cmd += ["-c", os.path.join(CSRC, src), "-o", obj]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src}:\n{res.stdout}\n{res.stderr}")
    if verbose and (res.stdout or res.stderr):
        print(res.stdout, res.stderr)
    return obj


def build_variant(name: str, defines: list[str]) -> str:
    """Tuning aid: libpcs_b200 with extra -D flags, written to csrc/build/variants/<name>.so (select it with
    PCS_B200_LIB=<path>)."""
    vdir = os.path.join(BUILD_DIR, "variants", name)
    os.makedirs(vdir, exist_ok=True)
    objs = []
    for src in sources():
        obj = os.path.join(vdir, src[:-3] + ".o")
        cmd = [nvcc(), *ARCH_FLAGS, *COMMON_FLAGS, *defines]
        if src in NO_FMAD:
            cmd.append("-fmad=false")
        subprocess.run(cmd + ["-c", os.path.join(CSRC, src), "-o", obj], check=True)
        objs.append(obj)
    out = os.path.join(BUILD_DIR, "variants", name + ".so")
    subprocess.run([nvcc(), *ARCH_FLAGS, "-shared", "-o", out, *objs], check=True)
    return out


def build(force: bool = False, verbose: bool = False, extra_flags: list[str] | None = None) -> str:
    """Compile every csrc/*.cu for sm_100a and link libpcs_b200.so.  Returns the library path."""
    os.makedirs(BUILD_DIR, exist_ok=True)
    extra = list(extra_flags or [])
    dep_t = _deps_mtime()
    todo, objs = [], []
    for src in sources():
        obj = os.path.join(BUILD_DIR, src[:-3] + ".o")
        objs.append(obj)
        if force or not os.path.exists(obj) or os.path.getmtime(obj) < max(dep_t, os.path.getmtime(os.path.join(CSRC, src))):
            todo.append(src)
    if todo:
        with ThreadPoolExecutor(max_workers=min(len(todo), os.cpu_count() or 4)) as ex:
            list(ex.map(lambda s: _compile(s, verbose, extra), todo))
    if todo or not os.path.exists(LIB_PATH):
        cmd = [nvcc(), *ARCH_FLAGS, "-shared", "-o", LIB_PATH, *objs, "-Xcompiler", "-fvisibility=hidden"]
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError(f"link failed:\n{res.stdout}\n{res.stderr}")
    return LIB_PATH


if __name__ == "__main__":
    import sys

    print(build(force="--force" in sys.argv, verbose=True, extra_flags=["-Xptxas", "-v"] if "-v" in sys.argv else None))
