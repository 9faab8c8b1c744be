"""ctypes binding of libpcs_b200.so (include/pcs_b200.h).

This is the only way Python reaches the kernels: every call goes through the C ABI.  There is no CPU path behind
it -- if the shared library is missing, or no CUDA device is present, the calls raise.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("PCS_B200_LIB") or os.path.join(_PKG, "libpcs_b200.so")  # env override: kernel variants

PCS_OK, PCS_ERR_BAD_ARG, PCS_ERR_CUDA, PCS_ERR_CAPACITY, PCS_ERR_COMM = 0, 1, 2, 3, 4
RETURN_INSIDE, RETURN_OUTSIDE, RETURN_COLORIZED = 0, 1, 2

POINT_DTYPE = np.dtype(
    [("x", "<f4"), ("y", "<f4"), ("z", "<f4"), ("data3", "<f4"), ("b", "u1"), ("g", "u1"), ("r", "u1"), ("a", "u1"),
     ("pad", "<u4", (3,))]
)
assert POINT_DTYPE.itemsize == 32
NOALLOC_MAX_POINTS = 1 << 23  # PCS_NOALLOC_MAX_POINTS
RECORD_DTYPE = np.dtype([("key48", "<u8"), ("counts", "<u8")])


class PcsError(RuntimeError):
    def __init__(self, status: int, message: str):
        super().__init__(f"pcs status {status}: {message}")
        self.status = status


class HsvParams(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("h_lo", "h_hi", "s_lo", "s_hi", "v_lo", "v_hi", "inverted", "morph_open")]


class AnnotateParams(C.Structure):
    _fields_ = [("hsv", HsvParams), ("lower", C.c_int32), ("upper", C.c_int32), ("limit_negative", C.c_int32),
                ("world_frame_output", C.c_int32)]


class Stats(C.Structure):
    _fields_ = [("kernel_launches", C.c_uint64), ("points_in", C.c_uint64), ("bytes_h2d", C.c_uint64),
                ("bytes_d2h", C.c_uint64)]


class MapStats(C.Structure):
    _fields_ = [("voxels", C.c_uint64), ("occupied", C.c_uint64), ("points_rejected", C.c_uint64),
                ("points_labelled", C.c_uint64), ("table_slots", C.c_uint64), ("load_factor", C.c_double)]


_vp, _sz, _i, _u64, _dbl = C.c_void_p, C.c_size_t, C.c_int, C.c_uint64, C.c_double
_pp = C.POINTER(C.c_void_p)
_pu64 = C.POINTER(C.c_uint64)

# name -> (restype, argtypes); mirrors include/pcs_b200.h declaration by declaration
SIGNATURES = {
    "pcs_version": (_i, []),
    "pcs_last_error": (C.c_char_p, []),
    "pcs_ctx_create": (_i, [_i, _vp, _pp]),
    "pcs_ctx_destroy": (_i, [_vp]),
    "pcs_ctx_sync": (_i, [_vp]),
    "pcs_get_stats": (_i, [_vp, C.POINTER(Stats)]),
    "pcs_hsv_default_params": (_i, [C.POINTER(HsvParams)]),
    "pcs_annotate_default_params": (_i, [C.POINTER(AnnotateParams)]),
    "pcs_host_alloc": (_vp, [_sz]),
    "pcs_host_free": (None, [_vp]),
    "pcs_dev_alloc": (_i, [_vp, _sz, _pp]),
    "pcs_dev_free": (_i, [_vp, _vp]),
    "pcs_memcpy_h2d": (_i, [_vp, _vp, _vp, _sz]),
    "pcs_memcpy_d2h": (_i, [_vp, _vp, _vp, _sz]),
    "pcs_memset_dev": (_i, [_vp, _vp, _i, _sz]),
    "pcs_event_create": (_i, [_vp, _pp]),
    "pcs_event_record": (_i, [_vp, _vp]),
    "pcs_stream_wait_event": (_i, [_vp, _vp]),
    "pcs_event_elapsed_ms": (_i, [_vp, _vp, C.POINTER(C.c_float)]),
    "pcs_event_destroy": (_i, [_vp]),
    "pcs_hsv_mask": (_i, [_vp, _vp, _i, _i, _sz, C.POINTER(HsvParams), _vp, _sz]),
    "pcs_hsv_mask_dev": (_i, [_vp, _vp, _i, _i, _sz, C.POINTER(HsvParams), _vp, _sz]),
    "pcs_hsv_annotation_image": (_i, [_vp, _vp, _i, _i, _sz, C.POINTER(HsvParams), _vp]),
    "pcs_histogram_backprojection": (_i, [_vp, _vp, _i, _i, _sz, _vp, _i, _i, _i, _i, _vp]),
    "pcs_cloud_to_image": (_i, [_vp, _vp, _i, _i, _vp, _vp]),
    "pcs_image_to_cloud": (_i, [_vp, _vp, _vp, _i, _i, _vp]),
    "pcs_color_passthrough": (_i, [_vp, _vp, _u64, _i, _i, _i, _vp, _pu64]),
    "pcs_color_passthrough_noalloc_dev": (_i, [_vp, _vp, _u64, _i, _i, _i, _vp, _vp]),
    "pcs_transform_cloud": (_i, [_vp, _vp, _vp, _u64, _vp]),
    "pcs_annotate_frames": (_i, [_vp, _vp, _vp, _i, _i, _i, _vp, C.POINTER(AnnotateParams), _vp, _vp]),
    "pcs_annotate_frames_dev": (_i, [_vp, _vp, _vp, _i, _i, _i, _vp, C.POINTER(AnnotateParams), _vp, _vp]),
    "pcs_labelled_points": (_i, [_vp, _vp, _i, _i, _i, _vp, C.POINTER(AnnotateParams), _dbl, _u64, _vp, _pu64]),
    "pcs_map_create": (_i, [_vp, _dbl, _u64, _pp]),
    "pcs_map_destroy": (_i, [_vp, _vp]),
    "pcs_map_clear": (_i, [_vp, _vp]),
    "pcs_map_integrate": (_i, [_vp, _vp, _vp, _u64, _vp, _i, _i, _i]),
    "pcs_map_integrate_dev": (_i, [_vp, _vp, _vp, _u64, _vp, _i, _i, _i]),
    "pcs_map_size": (_i, [_vp, _vp, _pu64, _pu64]),
    "pcs_map_get_stats": (_i, [_vp, _vp, C.POINTER(MapStats)]),
    "pcs_map_export": (_i, [_vp, _vp, _u64, _vp, _vp, _vp, _vp, _pu64]),
    "pcs_map_write_bt": (_i, [_vp, _vp, _vp, _u64, _pu64]),
    "pcs_map_leaf_centers": (_i, [_vp, _vp, _u64, _vp, _pu64]),
    "pcs_map_merge": (_i, [_vp, _vp, _vp, _vp, _vp, _u64]),
    "pcs_map_pack_by_owner_dev": (_i, [_vp, _vp, _i, _vp, _u64, _vp]),
    "pcs_map_merge_records_dev": (_i, [_vp, _vp, _vp, _u64]),
    "pcs_mix64": (_u64, [_u64]),
    "pcs_map_ipc_export": (_i, [_vp, _vp, _vp]),
    "pcs_map_ipc_open": (_i, [_vp, _vp, _dbl, _u64, _pp]),
    "pcs_map_push_to_owners_dev": (_i, [_vp, _vp, _pp, _i, _i]),
    "pcs_map_inbox_bytes": (_u64, [_u64, _i]),
    "pcs_ipc_export_dev": (_i, [_vp, _vp, _vp]),
    "pcs_ipc_open_dev": (_i, [_vp, _vp, _pp]),
    "pcs_ipc_close_dev": (_i, [_vp, _vp]),
    "pcs_map_push_mailbox_dev": (_i, [_vp, _vp, _vp, _i, _i, _pp, _u64, _i]),
    "pcs_map_merge_inbox_dev": (_i, [_vp, _vp, _vp, _i, _i, _u64]),
    "pcs_map_epoch_flag_dev": (_i, [_vp, _i, _pp]),
    "pcs_stream_write_u32_dev": (_i, [_vp, _vp, C.c_uint32]),
    "pcs_stream_wait_geq_u32_dev": (_i, [_vp, _vp, C.c_uint32]),
    "pcs_exchange_barrier_dev": (_i, [_vp, _pp, _pp, _i, C.c_uint32]),
    "pcs_mesh_mask": (_i, [_vp, _vp, _vp, _u64, _vp, _u64, _i, _u64, _vp, _vp, _vp, _vp, _pu64, _pu64]),
    "pcs_mesh_mask_dev": (_i, [_vp, _vp, _vp, _u64, _vp, _u64, _i, _u64, _vp, _vp, _vp, _vp, _vp]),
    "pcs_mesh_emit_dev": (_i, [_vp, _vp, _u64, _vp, _u64, _vp, _i, _u64, _vp, _vp, _vp, _vp]),
    "pcs_mesh_contacts_dev": (_i, [_vp, _vp, _vp, _u64, _vp, _u64, _vp]),
    "pcs_synth_frames_dev": (_i, [_vp, _u64, _i, _i, _i, _i, _vp, _vp]),
}

_lib = None


def load_library(path: str | None = None) -> C.CDLL:
    """dlopen libpcs_b200.so and type every exported symbol.  Raises if the library has not been built."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = path or LIB_PATH
    if not os.path.exists(p):
        raise RuntimeError(
            f"{p} not found: build it with `python -m point_cloud_segmentation_b200._build` "
            "(or __graft_entry__.build()). There is no CPU fallback.")
    L = C.CDLL(p)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(L, name)  # AttributeError here means the ABI and this table disagree
        fn.restype = res
        fn.argtypes = args
    if path is None:
        _lib = L
    return L


def _p(a):
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        return a.ctypes.data_as(C.c_void_p)
    return C.c_void_p(int(a))


def _check(status: int):
    if status != PCS_OK:
        raise PcsError(status, load_library().pcs_last_error().decode("utf-8", "replace"))


def _c(a, dtype):
    return np.ascontiguousarray(a, dtype=dtype)


def hsv_params(**kw) -> HsvParams:
    p = HsvParams()
    _check(load_library().pcs_hsv_default_params(C.byref(p)))
    for k, v in kw.items():
        setattr(p, k, int(v))
    return p


def annotate_params(hsv: HsvParams | None = None, **kw) -> AnnotateParams:
    p = AnnotateParams()
    _check(load_library().pcs_annotate_default_params(C.byref(p)))
    if hsv is not None:
        p.hsv = hsv
    for k, v in kw.items():
        setattr(p, k, int(v))
    return p


def mix64(x: int) -> int:
    return int(load_library().pcs_mix64(int(x)))


class PinnedBuffer:
    """cudaMallocHost memory exposed as a numpy array (for overlapped host<->device pipelines)."""

    def __init__(self, nbytes: int):
        self._lib = load_library()
        self.nbytes = int(nbytes)
        self.ptr = self._lib.pcs_host_alloc(self.nbytes)
        if not self.ptr:
            raise PcsError(PCS_ERR_CUDA, self._lib.pcs_last_error().decode())

    def array(self, dtype, count: int | None = None) -> np.ndarray:
        dtype = np.dtype(dtype)
        n = self.nbytes // dtype.itemsize if count is None else count
        buf = (C.c_uint8 * self.nbytes).from_address(self.ptr)
        return np.frombuffer(buf, dtype=dtype, count=n)

    def free(self):
        if self.ptr:
            self._lib.pcs_host_free(self.ptr)
            self.ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class VoxelMap:
    def __init__(self, ctx: "Context", resolution: float, capacity_voxels: int):
        self.ctx = ctx
        self.resolution = float(resolution)
        h = C.c_void_p()
        _check(ctx.lib.pcs_map_create(ctx.h, self.resolution, int(capacity_voxels), C.byref(h)))
        self.h = h

    def destroy(self):
        if self.h:
            self.ctx.lib.pcs_map_destroy(self.ctx.h, self.h)
            self.h = None

    def __del__(self):
        try:
            if self.ctx.h:
                self.destroy()
        except Exception:
            pass

    def clear(self):
        _check(self.ctx.lib.pcs_map_clear(self.ctx.h, self.h))

    def integrate(self, points: np.ndarray, pose=None, lower=0, upper=255, limit_negative=False):
        pts = _c(points, POINT_DTYPE)
        posep = None if pose is None else _c(pose, np.float32).reshape(16)
        _check(self.ctx.lib.pcs_map_integrate(self.ctx.h, self.h, _p(pts), pts.size, _p(posep), int(lower), int(upper),
                                              int(limit_negative)))

    def integrate_dev(self, points_dev: int, n: int, pose_dev: int | None = None, lower=0, upper=255,
                      limit_negative=False):
        _check(self.ctx.lib.pcs_map_integrate_dev(self.ctx.h, self.h, _p(points_dev), int(n), _p(pose_dev), int(lower),
                                                  int(upper), int(limit_negative)))

    def size(self):
        nv, no = C.c_uint64(0), C.c_uint64(0)
        _check(self.ctx.lib.pcs_map_size(self.ctx.h, self.h, C.byref(nv), C.byref(no)))
        return int(nv.value), int(no.value)

    def stats(self) -> dict:
        st = MapStats()
        _check(self.ctx.lib.pcs_map_get_stats(self.ctx.h, self.h, C.byref(st)))
        return {f: getattr(st, f) for f, _ in MapStats._fields_}

    def export(self):
        """(key48, count0, count1, logodds) in Morton order."""
        n = C.c_uint64(0)
        _check(self.ctx.lib.pcs_map_export(self.ctx.h, self.h, 0, None, None, None, None, C.byref(n)))
        m = int(n.value)
        k = np.empty(m, np.uint64)
        c0 = np.empty(m, np.uint32)
        c1 = np.empty(m, np.uint32)
        lo = np.empty(m, np.float32)
        if m:
            _check(self.ctx.lib.pcs_map_export(self.ctx.h, self.h, m, _p(k), _p(c0), _p(c1), _p(lo), C.byref(n)))
        return k, c0, c1, lo

    def write_bt(self) -> bytes:
        """Occupied voxels as an octomap binary tree (.bt) file image."""
        n = C.c_uint64(0)
        _check(self.ctx.lib.pcs_map_write_bt(self.ctx.h, self.h, None, 0, C.byref(n)))
        buf = np.empty(int(n.value), np.uint8)
        _check(self.ctx.lib.pcs_map_write_bt(self.ctx.h, self.h, _p(buf), buf.size, C.byref(n)))
        return buf.tobytes()

    def leaf_centers(self) -> np.ndarray:
        n = C.c_uint64(0)
        _check(self.ctx.lib.pcs_map_leaf_centers(self.ctx.h, self.h, 0, None, C.byref(n)))
        pts = np.empty(int(n.value), POINT_DTYPE)
        if pts.size:
            _check(self.ctx.lib.pcs_map_leaf_centers(self.ctx.h, self.h, pts.size, _p(pts), C.byref(n)))
        return pts

    def merge(self, key48, count0, count1):
        k = _c(key48, np.uint64)
        c0 = None if count0 is None else _c(count0, np.uint32)
        c1 = None if count1 is None else _c(count1, np.uint32)
        _check(self.ctx.lib.pcs_map_merge(self.ctx.h, self.h, _p(k), _p(c0), _p(c1), k.size))

    def pack_by_owner_dev(self, n_owners: int, records_dev: int, cap_records: int) -> np.ndarray:
        counts = np.zeros(n_owners, np.uint64)
        _check(self.ctx.lib.pcs_map_pack_by_owner_dev(self.ctx.h, self.h, int(n_owners), _p(records_dev),
                                                      int(cap_records), _p(counts)))
        return counts

    def ipc_export(self) -> bytes:
        """128-byte CUDA IPC handle pair of this map, to be opened by peer processes with VoxelMap.ipc_open."""
        buf = (C.c_uint8 * 128)()
        _check(self.ctx.lib.pcs_map_ipc_export(self.ctx.h, self.h, buf))
        return bytes(buf)

    @classmethod
    def ipc_open(cls, ctx: "Context", handle: bytes, resolution: float, capacity_voxels: int) -> "VoxelMap":
        self = cls.__new__(cls)
        self.ctx, self.resolution = ctx, float(resolution)
        h = C.c_void_p()
        buf = (C.c_uint8 * 128).from_buffer_copy(handle)
        _check(ctx.lib.pcs_map_ipc_open(ctx.h, buf, self.resolution, int(capacity_voxels), C.byref(h)))
        self.h = h
        return self

    def push_to_owners_dev(self, owners: "list[VoxelMap]", clear_local: bool = False):
        arr = (C.c_void_p * len(owners))(*[o.h for o in owners])
        _check(self.ctx.lib.pcs_map_push_to_owners_dev(self.ctx.h, self.h, arr, len(owners), int(clear_local)))

    def epoch_flag_dev(self, index: int) -> int:
        """Device address of exchange epoch word `index` of this (local or IPC-opened) map."""
        p = C.c_void_p()
        _check(self.ctx.lib.pcs_map_epoch_flag_dev(self.h, int(index), C.byref(p)))
        return int(p.value)

    def merge_records_dev(self, records_dev: int, n: int):
        _check(self.ctx.lib.pcs_map_merge_records_dev(self.ctx.h, self.h, _p(records_dev), int(n)))


class Context:
    """One CUDA device + stream (pcs_ctx)."""

    def __init__(self, device: int = 0, cuda_stream: int | None = None):
        self.lib = load_library()
        h = C.c_void_p()
        _check(self.lib.pcs_ctx_create(int(device), _p(cuda_stream), C.byref(h)))
        self.h = h
        self.device = device
        self.cuda_stream = int(cuda_stream) if cuda_stream else None  # None: the library created its own stream

    def destroy(self):
        if self.h:
            self.lib.pcs_ctx_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.destroy()
        except Exception:
            pass

    def sync(self):
        _check(self.lib.pcs_ctx_sync(self.h))

    def stats(self) -> dict:
        s = Stats()
        _check(self.lib.pcs_get_stats(self.h, C.byref(s)))
        return {f: int(getattr(s, f)) for f, _ in Stats._fields_}

    # ---- raw device memory / events -------------------------------------------------------------------------
    def dev_alloc(self, nbytes: int) -> int:
        d = C.c_void_p()
        _check(self.lib.pcs_dev_alloc(self.h, int(nbytes), C.byref(d)))
        return int(d.value)

    def dev_free(self, dptr: int):
        _check(self.lib.pcs_dev_free(self.h, _p(dptr)))

    def h2d(self, dptr: int, arr: np.ndarray):
        arr = np.ascontiguousarray(arr)
        _check(self.lib.pcs_memcpy_h2d(self.h, _p(dptr), _p(arr), arr.nbytes))

    def d2h(self, arr: np.ndarray, dptr: int):
        assert arr.flags.c_contiguous
        _check(self.lib.pcs_memcpy_d2h(self.h, _p(arr), _p(dptr), arr.nbytes))

    def memset(self, dptr: int, value: int, nbytes: int):
        _check(self.lib.pcs_memset_dev(self.h, _p(dptr), int(value), int(nbytes)))

    def ipc_export(self, dptr: int) -> bytes:
        buf = (C.c_uint8 * 64)()
        _check(self.lib.pcs_ipc_export_dev(self.h, _p(dptr), buf))
        return bytes(buf)

    def ipc_open(self, handle: bytes) -> int:
        buf = (C.c_uint8 * 64).from_buffer_copy(handle)
        d = C.c_void_p()
        _check(self.lib.pcs_ipc_open_dev(self.h, buf, C.byref(d)))
        return int(d.value)

    def ipc_close(self, dptr: int):
        _check(self.lib.pcs_ipc_close_dev(self.h, _p(dptr)))

    def push_mailbox(self, local: "VoxelMap", own: "VoxelMap", rank: int, peer_inboxes: "list[int]", seg_capacity: int,
                     clear_local: bool):
        arr = (C.c_void_p * len(peer_inboxes))(*[(p or None) for p in peer_inboxes])
        _check(self.lib.pcs_map_push_mailbox_dev(self.h, local.h, own.h, int(rank), len(peer_inboxes), arr, int(seg_capacity),
                                                 int(clear_local)))

    def merge_inbox(self, own: "VoxelMap", inbox: int, rank: int, n_owners: int, seg_capacity: int):
        _check(self.lib.pcs_map_merge_inbox_dev(self.h, own.h, _p(inbox), int(rank), int(n_owners), int(seg_capacity)))

    def stream_write_u32(self, flag_dev: int, value: int):
        _check(self.lib.pcs_stream_write_u32_dev(self.h, _p(flag_dev), int(value) & 0xFFFFFFFF))

    def stream_wait_geq_u32(self, flag_dev: int, value: int):
        _check(self.lib.pcs_stream_wait_geq_u32_dev(self.h, _p(flag_dev), int(value) & 0xFFFFFFFF))

    def exchange_barrier(self, their_flags: "list[int]", my_flags: "list[int]", epoch: int):
        n = len(their_flags)
        a = (C.c_void_p * max(n, 1))(*their_flags)
        b = (C.c_void_p * max(n, 1))(*my_flags)
        _check(self.lib.pcs_exchange_barrier_dev(self.h, a, b, n, int(epoch) & 0xFFFFFFFF))

    def event(self) -> int:
        e = C.c_void_p()
        _check(self.lib.pcs_event_create(self.h, C.byref(e)))
        return int(e.value)

    def record(self, ev: int):
        _check(self.lib.pcs_event_record(self.h, _p(ev)))

    def wait_event(self, ev: int):
        """Orders this context's stream after `ev` (an event recorded by any context of this process)."""
        _check(self.lib.pcs_stream_wait_event(self.h, _p(ev)))

    def elapsed_ms(self, start: int, stop: int) -> float:
        ms = C.c_float(0)
        _check(self.lib.pcs_event_elapsed_ms(_p(start), _p(stop), C.byref(ms)))
        return float(ms.value)

    def event_destroy(self, ev: int):
        self.lib.pcs_event_destroy(_p(ev))

    # ---- stage 1 -----------------------------------------------------------------------------------------------
    def hsv_mask(self, bgr: np.ndarray, params: HsvParams | None = None) -> np.ndarray:
        """bgr: (h, w, 3) uint8, rows may be strided (bgr.strides[0] is honoured)."""
        assert bgr.dtype == np.uint8 and bgr.ndim == 3 and bgr.shape[2] == 3 and bgr.strides[1:] == (3, 1)
        h, w, _ = bgr.shape
        mask = np.empty((h, w), np.uint8)
        p = params or hsv_params()
        _check(self.lib.pcs_hsv_mask(self.h, _p(bgr), w, h, bgr.strides[0], C.byref(p), _p(mask), mask.strides[0]))
        return mask

    def hsv_mask_into(self, bgr: np.ndarray, mask: np.ndarray, params: HsvParams | None = None):
        h, w, _ = bgr.shape
        p = params or hsv_params()
        _check(self.lib.pcs_hsv_mask(self.h, _p(bgr), w, h, bgr.strides[0], C.byref(p), _p(mask), mask.strides[0]))

    def hsv_annotation_image(self, bgr: np.ndarray, params: HsvParams | None = None) -> np.ndarray:
        assert bgr.dtype == np.uint8 and bgr.ndim == 3 and bgr.shape[2] == 3 and bgr.strides[1:] == (3, 1)
        h, w, _ = bgr.shape
        out = np.empty((h, w, 3), np.uint8)
        p = params or hsv_params()
        _check(self.lib.pcs_hsv_annotation_image(self.h, _p(bgr), w, h, bgr.strides[0], C.byref(p), _p(out)))
        return out

    def histogram_backprojection(self, bgr: np.ndarray, hist: np.ndarray, threshold_min: int = 30,
                                 threshold_value: int = 150) -> np.ndarray:
        """HistogramBackprojection.annotate_image: (h, w, 3) BGR + (h_bins, s_bins) float32 histogram -> (h, w, 3) mask."""
        assert bgr.dtype == np.uint8 and bgr.ndim == 3 and bgr.shape[2] == 3 and bgr.strides[1:] == (3, 1)
        hist = _c(hist, np.float32)
        h, w, _ = bgr.shape
        out = np.empty((h, w, 3), np.uint8)
        _check(self.lib.pcs_histogram_backprojection(self.h, _p(bgr), w, h, bgr.strides[0], _p(hist), hist.shape[0],
                                                     hist.shape[1], int(threshold_min), int(threshold_value), _p(out)))
        return out

    # ---- stage 2 -----------------------------------------------------------------------------------------------
    def cloud_to_image(self, cloud: np.ndarray, w: int, h: int):
        cloud = _c(cloud, POINT_DTYPE)
        pos = np.empty((h, w, 3), np.float64)
        bgr = np.empty((h, w, 3), np.uint8)
        _check(self.lib.pcs_cloud_to_image(self.h, _p(cloud), w, h, _p(pos), _p(bgr)))
        return pos, bgr

    def image_to_cloud(self, annotation_bgr: np.ndarray, pos: np.ndarray) -> np.ndarray:
        ann = _c(annotation_bgr, np.uint8)
        pos = _c(pos, np.float64)
        h, w, _ = ann.shape
        out = np.empty(h * w, POINT_DTYPE)
        _check(self.lib.pcs_image_to_cloud(self.h, _p(ann), _p(pos), w, h, _p(out)))
        return out

    def color_passthrough(self, cloud: np.ndarray, lower: int, upper: int, limit_negative: bool = False) -> np.ndarray:
        cloud = _c(cloud, POINT_DTYPE)
        out = np.empty(max(cloud.size, 1), POINT_DTYPE)
        n = C.c_uint64(0)
        _check(self.lib.pcs_color_passthrough(self.h, _p(cloud), cloud.size, int(lower), int(upper), int(limit_negative),
                                              _p(out), C.byref(n)))
        return out[: int(n.value)].copy()

    def color_passthrough_noalloc_dev(self, in_dev: int, n: int, lower: int, upper: int, limit_negative: bool,
                                      out_dev: int, n_out_dev: int):
        """Allocation-free filter on caller-owned device buffers (at most NOALLOC_MAX_POINTS points); asynchronous."""
        _check(self.lib.pcs_color_passthrough_noalloc_dev(self.h, _p(in_dev), int(n), int(lower), int(upper),
                                                          int(limit_negative), _p(out_dev), _p(n_out_dev)))

    def transform_cloud(self, pose: np.ndarray, cloud: np.ndarray) -> np.ndarray:
        cloud = _c(cloud, POINT_DTYPE)
        pose = _c(pose, np.float32).reshape(16)
        out = np.empty_like(cloud)
        _check(self.lib.pcs_transform_cloud(self.h, _p(pose), _p(cloud), cloud.size, _p(out)))
        return out

    def annotate_frames(self, clouds: np.ndarray, n_frames: int, w: int, h: int, poses=None,
                        params: AnnotateParams | None = None, vmap: VoxelMap | None = None, want_annotated=True,
                        want_mask=True, annotated_out: np.ndarray | None = None, mask_out: np.ndarray | None = None):
        clouds = _c(clouds, POINT_DTYPE)
        assert clouds.size == n_frames * w * h
        p = params or annotate_params()
        posep = None if poses is None else _c(poses, np.float32).reshape(-1)
        ann = annotated_out if annotated_out is not None else (np.empty(clouds.size, POINT_DTYPE) if want_annotated else None)
        mask = mask_out if mask_out is not None else (np.empty((n_frames, h, w), np.uint8) if want_mask else None)
        _check(self.lib.pcs_annotate_frames(self.h, vmap.h if vmap else None, _p(clouds), n_frames, w, h, _p(posep),
                                            C.byref(p), _p(ann), _p(mask)))
        return ann, mask

    def labelled_points(self, clouds: np.ndarray, n_frames: int, w: int, h: int, poses=None,
                        params: AnnotateParams | None = None, resolution: float = 0.0) -> np.ndarray:
        """Order-preserving list of the finite, in-range points whose annotation passes the label filter (world frame)."""
        clouds = _c(clouds, POINT_DTYPE)
        p = params or annotate_params()
        posep = None if poses is None else _c(poses, np.float32).reshape(-1)
        n = C.c_uint64(0)
        out = np.empty(max(clouds.size, 1), POINT_DTYPE)
        _check(self.lib.pcs_labelled_points(self.h, _p(clouds), n_frames, w, h, _p(posep), C.byref(p), float(resolution),
                                            out.size, _p(out), C.byref(n)))
        return out[: int(n.value)].copy()

    def annotate_frames_dev(self, clouds_dev: int, n_frames: int, w: int, h: int, poses_dev: int | None,
                            params: AnnotateParams, vmap: VoxelMap | None, annotated_dev: int | None,
                            mask_dev: int | None):
        _check(self.lib.pcs_annotate_frames_dev(self.h, vmap.h if vmap else None, _p(clouds_dev), n_frames, w, h,
                                                _p(poses_dev), C.byref(params), _p(annotated_dev), _p(mask_dev)))

    # ---- stage 3 / 4 ---------------------------------------------------------------------------------------------
    def create_map(self, resolution: float, capacity_voxels: int) -> VoxelMap:
        return VoxelMap(self, resolution, capacity_voxels)

    def mesh_mask(self, vmap: VoxelMap, verts: np.ndarray, faces4: np.ndarray, mask_type: int):
        """Returns (out_verts (n,3), out_faces4 (m,4), out_rgb (n,3), tri_flags (nt,))."""
        verts = _c(verts, np.float64).reshape(-1, 3)
        faces4 = _c(faces4, np.int32).reshape(-1, 4)
        nv, nt = verts.shape[0], faces4.shape[0]
        flags = np.zeros(nt, np.uint8)
        nov, nom = C.c_uint64(0), C.c_uint64(0)
        # size query, then the real call (two-call pattern)
        _check(self.lib.pcs_mesh_mask(self.h, vmap.h, _p(verts), nv, _p(faces4), nt, int(mask_type), 0, None, None, None,
                                      None, C.byref(nov), C.byref(nom)))
        n_v, n_t = int(nov.value), int(nom.value)
        ov = np.empty((max(n_v, 1), 3), np.float64)
        of = np.empty((max(n_t, 1), 4), np.int32)
        rgb = np.empty((max(n_v, 1), 3), np.uint8)
        _check(self.lib.pcs_mesh_mask(self.h, vmap.h, _p(verts), nv, _p(faces4), nt, int(mask_type), max(n_t, 1), _p(ov),
                                      _p(of), _p(rgb), _p(flags), C.byref(nov), C.byref(nom)))
        return ov[:n_v], of[:n_t], rgb[:n_v], flags

    def mesh_mask_dev(self, vmap: VoxelMap, verts_dev: int, nv: int, faces4_dev: int, nt: int, mask_type: int, cap_tris: int,
                      out_verts_dev: int | None, out_faces4_dev: int | None, out_rgb_dev: int | None, flags_dev: int,
                      n_out_dev: int | None):
        """Device-resident maskMesh: flags + compacted output (or colours), asynchronous on the context's stream."""
        _check(self.lib.pcs_mesh_mask_dev(self.h, vmap.h, _p(verts_dev), int(nv), _p(faces4_dev), int(nt), int(mask_type),
                                          int(cap_tris), _p(out_verts_dev), _p(out_faces4_dev), _p(out_rgb_dev), _p(flags_dev),
                                          _p(n_out_dev)))

    def mesh_emit_dev(self, verts_dev: int, nv: int, faces4_dev: int, nt: int, flags_dev: int, mask_type: int, cap_tris: int,
                      out_verts_dev: int | None, out_faces4_dev: int | None, out_rgb_dev: int | None, n_out_dev: int):
        """Output assembly from existing contact flags (multi-GPU mask: flags gathered from all ranks)."""
        _check(self.lib.pcs_mesh_emit_dev(self.h, _p(verts_dev), int(nv), _p(faces4_dev), int(nt), _p(flags_dev), int(mask_type),
                                          int(cap_tris), _p(out_verts_dev), _p(out_faces4_dev), _p(out_rgb_dev), _p(n_out_dev)))

    def mesh_contacts_dev(self, vmap: VoxelMap, verts_dev: int, nv: int, faces4_dev: int, nt: int, flags_dev: int):
        _check(self.lib.pcs_mesh_contacts_dev(self.h, vmap.h, _p(verts_dev), int(nv), _p(faces4_dev), int(nt),
                                              _p(flags_dev)))

    # ---- synthetic input ---------------------------------------------------------------------------------------
    def synth_frames_dev(self, seed: int, first_frame: int, n_frames: int, w: int, h: int, poses_dev: int, out_dev: int):
        _check(self.lib.pcs_synth_frames_dev(self.h, int(seed), int(first_frame), int(n_frames), w, h, _p(poses_dev),
                                             _p(out_dev)))
