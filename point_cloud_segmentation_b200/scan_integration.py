"""Python mirror of pcs_scan_integration (colorPassthrough, OctomapMeshMask) over the C ABI.

Names, argument meaning and defaults follow
pcs_scan_integration/include/pcs_scan_integration/octomap_mesh_masking.h:50-143; meshes are (vertices (n,3) float64,
faces4 (m,4) int32 = [3, i0, i1, i2]) pairs, the layout of tesseract_geometry::Mesh.
"""
from __future__ import annotations

import enum

import numpy as np

from . import capi
from .detection import default_context


def colorPassthrough(input_cloud: np.ndarray, lower_limit: int, upper_limit: int, limit_negative: bool = False,
                     context: capi.Context | None = None) -> np.ndarray:
    """Points whose r, g and b all lie in [lower_limit, upper_limit] (limit_negative: all outside), input order kept."""
    return (context or default_context()).color_passthrough(input_cloud, lower_limit, upper_limit, limit_negative)


class OctomapMeshMask:
    class MaskType(enum.IntEnum):
        RETURN_INSIDE = 0
        RETURN_OUTSIDE = 1
        RETURN_COLORIZED = 2

    def __init__(self, context: capi.Context | None = None):
        self._ctx = context or default_context()
        self._input_mesh = None
        self._masked_mesh = None
        self._colors = None
        self._octree = None

    # -- mesh ---------------------------------------------------------------------------------------------------
    def setInputMesh(self, vertices: np.ndarray, faces4: np.ndarray):
        self._input_mesh = (np.ascontiguousarray(vertices, np.float64).reshape(-1, 3),
                            np.ascontiguousarray(faces4, np.int32).reshape(-1, 4))

    def getInputMesh(self):
        return self._input_mesh

    # -- octree -------------------------------------------------------------------------------------------------
    def setOctree(self, point_cloud, resolution: float | None = None, lower_limit: int = 0, upper_limit: int = 255,
                  limit_negative: bool = False):
        """setOctree(cloud, resolution, lower, upper, limit_negative) or setOctree(voxel_map)."""
        if isinstance(point_cloud, capi.VoxelMap) or point_cloud is None:
            self._octree = point_cloud
            return
        cloud = np.ascontiguousarray(point_cloud, capi.POINT_DTYPE)
        vmap = self._ctx.create_map(float(resolution), 2 * cloud.size + 1024)
        vmap.integrate(cloud, None, lower_limit, upper_limit, limit_negative)
        self._octree = vmap

    def getOctree(self):
        return self._octree

    # -- masking ------------------------------------------------------------------------------------------------
    def maskMesh(self, mask_type) -> bool:
        if self._octree is None or self._input_mesh is None:
            return False
        try:
            v, f, rgb, _ = self._ctx.mesh_mask(self._octree, self._input_mesh[0], self._input_mesh[1], int(mask_type))
        except capi.PcsError:
            return False
        self._masked_mesh, self._colors = (v.copy(), f.copy()), rgb.copy()
        return True

    def getMaskedMesh(self):
        return self._masked_mesh

    def saveMaskedMesh(self, filepath: str) -> bool:
        """ASCII PLY with per-vertex colours, the layout of tesseract's writeSimplePlyFile."""
        if self._masked_mesh is None:
            return False
        v, f = self._masked_mesh
        try:
            with open(filepath, "w") as fh:
                fh.write("ply\nformat ascii 1.0\ncomment made by tesseract\n")
                fh.write(f"element vertex {v.shape[0]}\nproperty float x\nproperty float y\nproperty float z\n")
                fh.write("property uchar red\nproperty uchar green\nproperty uchar blue\n")
                fh.write(f"element face {f.shape[0]}\nproperty list uchar int vertex_indices\nend_header\n")
                for p, c in zip(v, self._colors):
                    fh.write("%.7f %.7f %.7f %d %d %d\n" % (p[0], p[1], p[2], c[0], c[1], c[2]))
                for t in f:
                    fh.write("%d %d %d %d\n" % tuple(t))
            return True
        except OSError:
            return False
