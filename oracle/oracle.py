"""ctypes loader for the CPU oracle (oracle/pcs_oracle.h).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
``--impl reference`` legs.  The product package (point_cloud_segmentation_b200) never imports this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_build", "libpcs_oracle.so")

POINT_DTYPE = np.dtype(
    [("x", "<f4"), ("y", "<f4"), ("z", "<f4"), ("w", "<f4"), ("b", "u1"), ("g", "u1"), ("r", "u1"), ("a", "u1"),
     ("pad", "<u4", (3,))]
)
assert POINT_DTYPE.itemsize == 32


class HsvParams(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("h_lo", "h_hi", "s_lo", "s_hi", "v_lo", "v_hi", "inverted", "morph_open")]


class BaselineResult(C.Structure):
    _fields_ = [("seconds_annotate", C.c_double), ("seconds_integrate", C.c_double), ("points", C.c_uint64),
                ("inserted", C.c_uint64), ("leaves", C.c_uint64)]


def build(force: bool = False) -> str:
    """Compile oracle/*.c with gcc (make) into oracle/_build/libpcs_oracle.so."""
    srcs = [os.path.join(_HERE, f) for f in os.listdir(_HERE) if f.endswith((".c", ".h")) or f == "Makefile"]
    stale = force or not os.path.exists(_LIB_PATH) or any(
        os.path.getmtime(s) > os.path.getmtime(_LIB_PATH) for s in srcs)
    if stale:
        subprocess.run(["make", "-C", _HERE, "-s"], check=True)
    return _LIB_PATH


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB_PATH)
        vp, sz, i32, u64, dbl = C.c_void_p, C.c_size_t, C.c_int, C.c_uint64, C.c_double
        sig = {
            "orc_hsv_default_params": (None, [C.POINTER(HsvParams)]),
            "orc_hsv_tables": (None, [vp, vp]),
            "orc_bgr2hsv": (None, [vp, i32, i32, sz, vp, sz]),
            "orc_hsv_mask": (None, [vp, i32, i32, sz, C.POINTER(HsvParams), vp, sz]),
            "orc_morph_open5": (None, [vp, i32, i32, vp]),
            "orc_hsv_exhaustive_hash": (u64, [C.POINTER(HsvParams), C.POINTER(u64)]),
            "orc_histogram_backprojection": (None, [vp, i32, i32, sz, vp, i32, i32, i32, i32, vp]),
            "orc_cloud_to_image": (None, [vp, i32, i32, vp, vp]),
            "orc_image_to_cloud": (None, [vp, vp, i32, i32, vp]),
            "orc_color_passthrough": (sz, [vp, sz, i32, i32, i32, vp]),
            "orc_transform_points": (None, [vp, vp, sz, vp]),
            "orc_annotate_frame": (None, [vp, i32, i32, C.POINTER(HsvParams), vp, vp]),
            "orc_octree_create": (vp, [dbl]),
            "orc_octree_destroy": (None, [vp]),
            "orc_octree_resolution": (dbl, [vp]),
            "orc_coord_to_key": (i32, [dbl, dbl, C.POINTER(C.c_uint16)]),
            "orc_octree_insert_point": (i32, [vp, C.c_float, C.c_float, C.c_float]),
            "orc_octree_insert_key": (None, [vp, C.c_uint16, C.c_uint16, C.c_uint16]),
            "orc_octree_insert_cloud": (None, [vp, vp, sz]),
            "orc_octree_update_inner_occupancy": (None, [vp]),
            "orc_octree_prune_tesseract": (None, [vp]),
            "orc_octree_prune_octomap": (None, [vp]),
            "orc_octree_to_max_likelihood": (None, [vp]),
            "orc_octree_num_nodes": (sz, [vp]),
            "orc_octree_num_leaves": (sz, [vp]),
            "orc_octree_leaves": (sz, [vp, vp, vp, vp, vp, vp, sz]),
            "orc_octree_search": (i32, [vp, C.c_uint16, C.c_uint16, C.c_uint16, C.POINTER(C.c_float),
                                        C.POINTER(C.c_int)]),
            "orc_octree_read_bt": (vp, [vp, sz]),
            "orc_octree_write_bt": (sz, [vp, vp, sz]),
            "orc_logodds_after_hits": (C.c_float, [C.c_uint32]),
            "orc_morton48": (u64, [C.c_uint16, C.c_uint16, C.c_uint16]),
            "orc_count_voxels": (sz, [vp, sz, vp, dbl, i32, i32, i32, vp, vp, vp, sz]),
            "orc_tri_box_overlap": (i32, [vp, vp, vp, vp, vp]),
            "orc_mesh_contacts_tree": (u64, [vp, vp, sz, vp, sz, vp]),
            "orc_mesh_contacts_keys": (u64, [vp, sz, dbl, vp, sz, vp, sz, vp]),
            "orc_mesh_contacts_brute": (u64, [vp, vp, sz, vp, vp, sz, vp]),
            "orc_mask_mesh_emit": (sz, [vp, sz, vp, sz, vp, i32, vp, vp, vp, C.POINTER(sz)]),
            "orc_load_stl_binary": (i32, [vp, sz, vp, sz, vp, sz, C.POINTER(sz), C.POINTER(sz)]),
            "orc_load_ply_ascii": (i32, [vp, sz, vp, sz, vp, sz, C.POINTER(sz), C.POINTER(sz)]),
            "orc_write_ply_ascii": (sz, [vp, sz, vp, vp, sz, vp, sz]),
            "orc_splitmix64": (u64, [u64]),
            "orc_psin": (dbl, [dbl]),
            "orc_pcos": (dbl, [dbl]),
            "orc_synth_pose": (None, [i32, vp]),
            "orc_synth_frame": (None, [u64, i32, i32, i32, vp, vp]),
            "orc_baseline_annotate_integrate": (None, [vp, i32, i32, i32, vp, C.POINTER(HsvParams), dbl, i32, i32,
                                                       i32, i32, C.POINTER(BaselineResult)]),
        }
        for name, (res, args) in sig.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _c(a, dtype):
    return np.ascontiguousarray(a, dtype=dtype)


def default_params(**kw) -> HsvParams:
    p = HsvParams()
    lib().orc_hsv_default_params(C.byref(p))
    for k, v in kw.items():
        setattr(p, k, int(v))
    return p


# ---- stage 1 -------------------------------------------------------------------------------------------
def bgr2hsv(bgr: np.ndarray) -> np.ndarray:
    bgr = _c(bgr, np.uint8)
    h, w, _ = bgr.shape
    out = np.empty_like(bgr)
    lib().orc_bgr2hsv(_p(bgr), w, h, w * 3, _p(out), w * 3)
    return out


def hsv_mask(bgr: np.ndarray, params: HsvParams | None = None) -> np.ndarray:
    bgr = _c(bgr, np.uint8)
    h, w, _ = bgr.shape
    p = params or default_params()
    out = np.empty((h, w), np.uint8)
    lib().orc_hsv_mask(_p(bgr), w, h, w * 3, C.byref(p), _p(out), w)
    return out


def morph_open5(mask: np.ndarray) -> np.ndarray:
    mask = _c(mask, np.uint8)
    h, w = mask.shape
    out = np.empty_like(mask)
    lib().orc_morph_open5(_p(mask), w, h, _p(out))
    return out


def hsv_exhaustive_hash(params: HsvParams | None = None):
    p = params or default_params()
    n = C.c_uint64(0)
    hsh = lib().orc_hsv_exhaustive_hash(C.byref(p), C.byref(n))
    return int(hsh), int(n.value)


def histogram_backprojection(bgr: np.ndarray, hist: np.ndarray, thr_min: int = 30, thr_value: int = 150) -> np.ndarray:
    bgr = _c(bgr, np.uint8)
    hist = _c(hist, np.float32)
    h, w, _ = bgr.shape
    out = np.empty((h, w, 3), np.uint8)
    lib().orc_histogram_backprojection(_p(bgr), w, h, w * 3, _p(hist), hist.shape[0], hist.shape[1], thr_min, thr_value,
                                       _p(out))
    return out


# ---- stage 2 -------------------------------------------------------------------------------------------
def cloud_to_image(cloud: np.ndarray, w: int, h: int):
    cloud = _c(cloud, POINT_DTYPE)
    pos = np.empty((h, w, 3), np.float64)
    bgr = np.empty((h, w, 3), np.uint8)
    lib().orc_cloud_to_image(_p(cloud), w, h, _p(pos), _p(bgr))
    return pos, bgr


def image_to_cloud(annotation_bgr: np.ndarray, pos: np.ndarray) -> np.ndarray:
    ann = _c(annotation_bgr, np.uint8)
    pos = _c(pos, np.float64)
    h, w, _ = ann.shape
    out = np.empty(h * w, POINT_DTYPE)
    lib().orc_image_to_cloud(_p(ann), _p(pos), w, h, _p(out))
    return out


def color_passthrough(cloud: np.ndarray, lower: int, upper: int, limit_negative: bool = False) -> np.ndarray:
    cloud = _c(cloud, POINT_DTYPE)
    out = np.empty_like(cloud)
    n = lib().orc_color_passthrough(_p(cloud), cloud.size, lower, upper, int(limit_negative), _p(out))
    return out[:n].copy()


def transform_points(pose: np.ndarray, cloud: np.ndarray) -> np.ndarray:
    cloud = _c(cloud, POINT_DTYPE)
    pose = _c(pose, np.float32).reshape(16)
    out = np.empty_like(cloud)
    lib().orc_transform_points(_p(pose), _p(cloud), cloud.size, _p(out))
    return out


def annotate_frame(cloud: np.ndarray, w: int, h: int, params: HsvParams | None = None):
    cloud = _c(cloud, POINT_DTYPE)
    p = params or default_params()
    out = np.empty(w * h, POINT_DTYPE)
    mask = np.empty((h, w), np.uint8)
    lib().orc_annotate_frame(_p(cloud), w, h, C.byref(p), _p(out), _p(mask))
    return out, mask


# ---- stage 3 -------------------------------------------------------------------------------------------
class Octree:
    def __init__(self, resolution: float | None = None, handle=None):
        self._h = handle if handle is not None else lib().orc_octree_create(float(resolution))
        if not self._h:
            raise ValueError("octree creation failed")

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orc_octree_destroy(self._h)
            self._h = None

    @classmethod
    def read_bt(cls, data: bytes) -> "Octree":
        buf = np.frombuffer(data, np.uint8)
        h = lib().orc_octree_read_bt(_p(buf), buf.size)
        if not h:
            raise ValueError("bad .bt data")
        return cls(handle=h)

    def write_bt(self) -> bytes:
        n = lib().orc_octree_write_bt(self._h, None, 0)
        buf = np.empty(n, np.uint8)
        lib().orc_octree_write_bt(self._h, _p(buf), n)
        return buf.tobytes()

    @property
    def resolution(self) -> float:
        return lib().orc_octree_resolution(self._h)

    def insert_cloud(self, cloud: np.ndarray):
        cloud = _c(cloud, POINT_DTYPE)
        lib().orc_octree_insert_cloud(self._h, _p(cloud), cloud.size)

    def insert_key(self, kx, ky, kz):
        lib().orc_octree_insert_key(self._h, int(kx), int(ky), int(kz))

    def finalize_tesseract(self):
        """updateInnerOccupancy + tesseract prune, as tesseract_geometry::Octree(cloud, res, BOX, true)."""
        lib().orc_octree_update_inner_occupancy(self._h)
        lib().orc_octree_prune_tesseract(self._h)

    def num_nodes(self) -> int:
        return lib().orc_octree_num_nodes(self._h)

    def leaves(self):
        n = lib().orc_octree_num_leaves(self._h)
        kx = np.empty(n, np.uint16)
        ky = np.empty(n, np.uint16)
        kz = np.empty(n, np.uint16)
        d = np.empty(n, np.uint8)
        lo = np.empty(n, np.float32)
        lib().orc_octree_leaves(self._h, _p(kx), _p(ky), _p(kz), _p(d), _p(lo), n)
        return kx, ky, kz, d, lo

    def leaf_voxels(self, occupied_only: bool = True):
        """Expand every leaf to finest-level voxels; returns (key48 packed sorted, logodds)."""
        kx, ky, kz, d, lo = self.leaves()
        keys, vals = [], []
        for x, y, z, dd, v in zip(kx.astype(np.int64), ky.astype(np.int64), kz.astype(np.int64), d, lo):
            if occupied_only and v < 0:
                continue
            s = 1 << (16 - int(dd))
            r = np.arange(s, dtype=np.int64)
            gx, gy, gz = np.meshgrid(x + r, y + r, z + r, indexing="ij")
            k = (gz.ravel() << 32) | (gy.ravel() << 16) | gx.ravel()
            keys.append(k.astype(np.uint64))
            vals.append(np.full(k.size, v, np.float32))
        if not keys:
            return np.empty(0, np.uint64), np.empty(0, np.float32)
        keys = np.concatenate(keys)
        vals = np.concatenate(vals)
        o = np.argsort(keys)
        return keys[o], vals[o]

    def search(self, kx, ky, kz):
        lo = C.c_float(0)
        d = C.c_int(0)
        ok = lib().orc_octree_search(self._h, int(kx), int(ky), int(kz), C.byref(lo), C.byref(d))
        return (lo.value, d.value) if ok else None

    def mesh_contacts(self, verts: np.ndarray, faces4: np.ndarray) -> np.ndarray:
        verts = _c(verts, np.float64)
        faces4 = _c(faces4, np.int32)
        nt = faces4.size // 4
        pc = np.zeros(nt, np.uint32)
        lib().orc_mesh_contacts_tree(self._h, _p(verts), verts.size // 3, _p(faces4), nt, _p(pc))
        return pc


def logodds_after_hits(c: int) -> np.float32:
    return np.float32(lib().orc_logodds_after_hits(int(c)))


def morton48(key48: np.ndarray) -> np.ndarray:
    k = np.asarray(key48, np.uint64)

    def spread(v):
        v = v & np.uint64(0xFFFF)
        out = np.zeros_like(v)
        for i in range(16):
            out |= ((v >> np.uint64(i)) & np.uint64(1)) << np.uint64(3 * i)
        return out

    return spread(k) | (spread(k >> np.uint64(16)) << np.uint64(1)) | (spread(k >> np.uint64(32)) << np.uint64(2))


def count_voxels(points: np.ndarray, resolution: float, pose=None, lower=0, upper=255, limit_negative=False):
    """Exact (key48, count0, count1) records in Morton order: count1 = points passing colorPassthrough."""
    pts = _c(points, POINT_DTYPE)
    posep = None if pose is None else _c(pose, np.float32).reshape(16)
    cap = max(pts.size, 1)
    k = np.empty(cap, np.uint64)
    c0 = np.empty(cap, np.uint32)
    c1 = np.empty(cap, np.uint32)
    n = lib().orc_count_voxels(_p(pts), pts.size, _p(posep), float(resolution), lower, upper, int(limit_negative),
                               _p(k), _p(c0), _p(c1), cap)
    return k[:n].copy(), c0[:n].copy(), c1[:n].copy()


# ---- stage 4 -------------------------------------------------------------------------------------------
def tri_box_overlap(lo, hi, v0, v1, v2) -> bool:
    a = [_c(x, np.float64) for x in (lo, hi, v0, v1, v2)]
    return bool(lib().orc_tri_box_overlap(*[_p(x) for x in a]))


def mesh_contacts_keys(key48: np.ndarray, resolution: float, verts: np.ndarray, faces4: np.ndarray) -> np.ndarray:
    key48 = _c(key48, np.uint64)
    verts = _c(verts, np.float64)
    faces4 = _c(faces4, np.int32)
    nt = faces4.size // 4
    flags = np.zeros(nt, np.uint8)
    lib().orc_mesh_contacts_keys(_p(key48), key48.size, float(resolution), _p(verts), verts.size // 3, _p(faces4), nt,
                                 _p(flags))
    return flags


def mesh_contacts_brute(box_lo, box_hi, verts, faces4) -> np.ndarray:
    box_lo = _c(box_lo, np.float64)
    box_hi = _c(box_hi, np.float64)
    verts = _c(verts, np.float64)
    faces4 = _c(faces4, np.int32)
    nt = faces4.size // 4
    pc = np.zeros(nt, np.uint32)
    lib().orc_mesh_contacts_brute(_p(box_lo), _p(box_hi), box_lo.size // 3, _p(verts), _p(faces4), nt, _p(pc))
    return pc


def mask_mesh_emit(verts, faces4, tri_flags, mask_type: int):
    verts = _c(verts, np.float64).reshape(-1, 3)
    faces4 = _c(faces4, np.int32).reshape(-1, 4)
    flags = _c(tri_flags, np.uint8)
    nv, nt = verts.shape[0], faces4.shape[0]
    cap_v = max(nv, 3 * nt, 1)
    ov = np.empty((cap_v, 3), np.float64)
    of = np.empty((max(nt, 1), 4), np.int32)
    rgb = np.empty((cap_v, 3), np.uint8)
    nov = C.c_size_t(0)
    n = lib().orc_mask_mesh_emit(_p(verts), nv, _p(faces4), nt, _p(flags), mask_type, _p(ov), _p(of), _p(rgb),
                                 C.byref(nov))
    return ov[:nov.value].copy(), of[:n].copy(), rgb[:nov.value].copy()


def load_stl_binary(data: bytes):
    buf = np.frombuffer(data, np.uint8)
    nt_guess = max((buf.size - 84) // 50, 1)
    verts = np.empty((3 * nt_guess, 3), np.float64)
    faces = np.empty((nt_guess, 4), np.int32)
    nv, nt = C.c_size_t(0), C.c_size_t(0)
    rc = lib().orc_load_stl_binary(_p(buf), buf.size, _p(verts), verts.shape[0], _p(faces), faces.shape[0],
                                   C.byref(nv), C.byref(nt))
    if rc != 0:
        raise ValueError("bad STL")
    return verts[:nv.value].copy(), faces[:nt.value].copy()


def load_ply_ascii(text: bytes):
    buf = np.frombuffer(text, np.uint8)
    cap = max(buf.size // 4, 16)
    verts = np.empty((cap, 3), np.float64)
    faces = np.empty((cap, 4), np.int32)
    nv, nt = C.c_size_t(0), C.c_size_t(0)
    rc = lib().orc_load_ply_ascii(_p(buf), buf.size, _p(verts), cap, _p(faces), cap, C.byref(nv), C.byref(nt))
    if rc != 0:
        raise ValueError("bad PLY")
    return verts[:nv.value].copy(), faces[:nt.value].copy()


def write_ply_ascii(verts, rgb, faces4) -> bytes:
    verts = _c(verts, np.float64)
    faces4 = _c(faces4, np.int32)
    rgbp = None if rgb is None else _c(rgb, np.uint8)
    nv, nt = verts.size // 3, faces4.size // 4
    n = lib().orc_write_ply_ascii(_p(verts), nv, _p(rgbp), _p(faces4), nt, None, 0)
    buf = np.empty(n + 1, np.uint8)
    lib().orc_write_ply_ascii(_p(verts), nv, _p(rgbp), _p(faces4), nt, _p(buf), n + 1)
    return buf[:n].tobytes()


# ---- synthetic scan --------------------------------------------------------------------------------------
SEED = 20191002


def synth_pose(frame: int) -> np.ndarray:
    pose = np.empty(16, np.float32)
    lib().orc_synth_pose(int(frame), _p(pose))
    return pose


def synth_frame(frame: int, w: int, h: int, pose: np.ndarray | None = None, seed: int = SEED) -> np.ndarray:
    pose = synth_pose(frame) if pose is None else _c(pose, np.float32).reshape(16)
    out = np.empty(w * h, POINT_DTYPE)
    lib().orc_synth_frame(seed, int(frame), w, h, _p(pose), _p(out))
    return out


def baseline_annotate_integrate(frames: np.ndarray, n_frames: int, w: int, h: int, poses, resolution: float,
                                lower: int = 255, upper: int = 255, limit_negative: bool = False, threads: int = 1,
                                params: HsvParams | None = None) -> dict:
    frames = _c(frames, POINT_DTYPE)
    posesp = None if poses is None else _c(poses, np.float32).reshape(-1)
    p = params or default_params()
    res = BaselineResult()
    lib().orc_baseline_annotate_integrate(_p(frames), n_frames, w, h, _p(posesp), C.byref(p), float(resolution), lower,
                                          upper, int(limit_negative), threads, C.byref(res))
    return {f: getattr(res, f) for f, _ in BaselineResult._fields_}
