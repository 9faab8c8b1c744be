"""Shared helpers for the test-suite (input builders only; no product logic)."""
import numpy as np

SEED = 20191002


def faces4_of(tris: np.ndarray) -> np.ndarray:
    """(n,3) vertex indices -> tesseract layout [3, i0, i1, i2] per face."""
    tris = np.asarray(tris, np.int32).reshape(-1, 3)
    out = np.empty((tris.shape[0], 4), np.int32)
    out[:, 0] = 3
    out[:, 1:] = tris
    return out


def grid_mesh(nx: int, ny: int, x0: float, x1: float, y0: float, y1: float, zfun=None):
    """Height-field of nx x ny quads split into two triangles each (SURVEY 8d meshes)."""
    xs = np.linspace(x0, x1, nx + 1)
    ys = np.linspace(y0, y1, ny + 1)
    gx, gy = np.meshgrid(xs, ys, indexing="xy")
    gz = np.zeros_like(gx) if zfun is None else zfun(gx, gy)
    verts = np.stack([gx.ravel(), gy.ravel(), gz.ravel()], axis=1).astype(np.float64)
    idx = np.arange((nx + 1) * (ny + 1)).reshape(ny + 1, nx + 1)
    a, b, c, d = idx[:-1, :-1].ravel(), idx[:-1, 1:].ravel(), idx[1:, 1:].ravel(), idx[1:, :-1].ravel()
    tris = np.concatenate([np.stack([a, b, c], 1), np.stack([a, c, d], 1)], axis=0)
    # interleave the two triangles of each quad so that input order is spatially coherent
    tris = tris.reshape(2, -1, 3).transpose(1, 0, 2).reshape(-1, 3)
    return verts, faces4_of(tris)


def surface(x, y):
    return 0.02 * np.sin(8.0 * x) * np.cos(6.0 * y)


def ramp_cloud(point_dtype, n=255):
    """The reference's test cloud: 255 x 1, x = i/100, grey ramp (utils_unit.cpp:16-35,
    octomap_mesh_masking_unit.cpp:21-40)."""
    c = np.zeros(n, point_dtype)
    i = np.arange(n)
    c["x"] = (i.astype(np.float32) / np.float32(100.0)).astype(np.float32)
    c["r"] = c["g"] = c["b"] = i.astype(np.uint8)
    c["a"] = 255
    names = point_dtype.names
    c[names[3]] = 1.0
    return c


def random_cloud(point_dtype, n, rng, extent=1.0, nan_frac=0.05, inf_frac=0.01):
    c = np.zeros(n, point_dtype)
    for f in ("x", "y", "z"):
        c[f] = rng.uniform(-extent, extent, n).astype(np.float32)
    bad = rng.random(n) < nan_frac
    c["x"][bad] = np.nan
    bad2 = rng.random(n) < inf_frac
    c["y"][bad2] = np.inf
    bad3 = rng.random(n) < inf_frac
    c["z"][bad3] = -np.inf
    for f in ("r", "g", "b"):
        c[f] = rng.integers(0, 256, n, dtype=np.uint8)
    c["a"] = 255
    c[point_dtype.names[3]] = 1.0
    return c


def same_bytes(a: np.ndarray, b: np.ndarray) -> bool:
    return a.shape == b.shape and a.tobytes() == b.tobytes()


def key_boundary_cloud(point_dtype, res: float):
    """Points on the x axis at every voxel boundary k * res of the 16-bit key range (float32), one float32 ulp below and
    above each, and a few beyond both ends of the range (those must be culled)."""
    k = np.arange(-32768 - 2, 32768 + 3, dtype=np.float64)
    x = (k * res).astype(np.float32)
    xs = np.concatenate([x, np.nextafter(x, np.float32(-np.inf)), np.nextafter(x, np.float32(np.inf))])
    c = np.zeros(xs.size, point_dtype)
    c["x"] = xs
    c["r"] = c["g"] = c["b"] = 255
    c[point_dtype.names[3]] = 1.0
    return c


def expected_key_histogram(xs: np.ndarray, res: float):
    """Exact evaluation of octomap's coordToKeyChecked for float32 coordinates: floor(RN64(fl64(1 / res) * x)) + 32768,
    kept iff in [0, 65536).  Returns (sorted keys, counts)."""
    from fractions import Fraction
    import math

    inv = Fraction(1.0 / res)  # the double the library computes once per map
    hist = {}
    for x in xs.tolist():
        p = float(inv * Fraction(x))  # float(Fraction) rounds to nearest even: the single rounding of the double product
        k = math.floor(p) + 32768
        if 0 <= k < 65536:
            hist[k] = hist.get(k, 0) + 1
    keys = np.array(sorted(hist), np.uint64)
    return keys, np.array([hist[int(q)] for q in keys], np.uint64)
