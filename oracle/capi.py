"""ctypes front end of oracle/points_to_voxel.c (ORACLE -- test infrastructure).

``OracleVoxelGenerator`` mirrors the surface of ``spconv.utils.VoxelGenerator``
as the reference constructs it (second/builder/voxel_builder.py:23-27) so the
parity tests read like calls into the reference.
"""
import ctypes
import os
import subprocess

import numpy as np

OVERFLOW_BREAK = 0
OVERFLOW_CONTINUE = 1
_OVERFLOW = {"break": OVERFLOW_BREAK, "continue": OVERFLOW_CONTINUE}

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

_f32p = ctypes.POINTER(ctypes.c_float)
_i32p = ctypes.POINTER(ctypes.c_int32)
_i64p = ctypes.POINTER(ctypes.c_int64)


def library_path():
    return os.path.join(_HERE, "_build", "liboracle.so")


def build_library(force=False):
    """Compile oracle/points_to_voxel.c with gcc (recipe: oracle/Makefile)."""
    src = os.path.join(_HERE, "points_to_voxel.c")
    out = library_path()
    if force or not os.path.exists(out) or os.path.getmtime(out) < os.path.getmtime(src):
        subprocess.run(["make", "-C", _HERE, "-B"], check=True, capture_output=True)
    return out


def _lib():
    global _LIB
    if _LIB is None:
        if not os.path.exists(library_path()):
            build_library()
        _LIB = ctypes.CDLL(library_path())
        _LIB.orc_points_to_voxel.restype = ctypes.c_int
        _LIB.orc_points_to_voxel.argtypes = [
            _f32p, ctypes.c_int, ctypes.c_int, _f32p, _f32p, _i32p,
            ctypes.c_int, ctypes.c_int, ctypes.c_int, _i32p, _f32p, _i32p, _i32p]
        _LIB.orc_grid_size.restype = None
        _LIB.orc_grid_size.argtypes = [_f32p, _f32p, _i64p]
        _LIB.orc_transform_sweep.restype = None
        _LIB.orc_transform_sweep.argtypes = [
            _f32p, ctypes.c_int, ctypes.c_int, _f32p, _f32p, ctypes.c_float,
            ctypes.c_int, _f32p]
        _LIB.orc_mean_vfe.restype = None
        _LIB.orc_mean_vfe.argtypes = [
            _f32p, _i32p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, _f32p]
        _LIB.orc_scatter_bev.restype = None
        _LIB.orc_scatter_bev.argtypes = [
            _f32p, _i32p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int,
            ctypes.c_int, ctypes.c_int, _f32p]
        _LIB.orc_sparse_sum_for_anchors_mask.restype = None
        _LIB.orc_sparse_sum_for_anchors_mask.argtypes = [
            _i32p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int, _f32p]
        _LIB.orc_cumsum2d.restype = None
        _LIB.orc_cumsum2d.argtypes = [_f32p, ctypes.c_int, ctypes.c_int]
        _LIB.orc_fused_get_anchors_area.restype = None
        _LIB.orc_fused_get_anchors_area.argtypes = [
            _f32p, ctypes.c_int, ctypes.c_int, _f32p, ctypes.c_int, _f32p, _f32p, _i64p, _f32p]
    return _LIB


def _p(a, t):
    return a.ctypes.data_as(t)


def grid_size(voxel_size, point_cloud_range):
    """Appendix A.1: np.round((range[3:]-range[:3])/voxel_size) in fp32 -> int64 xyz."""
    vs = np.ascontiguousarray(voxel_size, dtype=np.float32)
    rg = np.ascontiguousarray(point_cloud_range, dtype=np.float32)
    out = np.zeros(3, dtype=np.int64)
    _lib().orc_grid_size(_p(vs, _f32p), _p(rg, _f32p), _p(out, _i64p))
    return out


class OracleVoxelGenerator:
    """spconv.utils.VoxelGenerator twin (Appendix A.1/A.2) on the C kernel.

    Constructor signature follows second/builder/voxel_builder.py:23-27.
    ``overflow`` selects what happens when voxel ``max_voxels`` would be created:
    "break" (spconv v1.0 / second.pytorch v1.5, default) or "continue".
    """

    def __init__(self, voxel_size, point_cloud_range, max_num_points,
                 max_voxels=20000, overflow="break"):
        self._voxel_size = np.array(voxel_size, dtype=np.float32)
        self._point_cloud_range = np.array(point_cloud_range, dtype=np.float32)
        self._grid_size = grid_size(self._voxel_size, self._point_cloud_range)
        self._max_num_points = int(max_num_points)
        self._max_voxels = int(max_voxels)
        self._overflow = _OVERFLOW[overflow]
        # coor_to_voxelidx, zyx, cached on the object like the reference does
        self._grid = None

    def _dense_grid(self):
        if self._grid is None:
            self._grid = np.full(tuple(self._grid_size[::-1]), -1, dtype=np.int32)
        return self._grid

    def generate(self, points, max_voxels=None):
        max_voxels = int(max_voxels or self._max_voxels)
        points = np.ascontiguousarray(points, dtype=np.float32)
        n, f = points.shape
        t = self._max_num_points
        voxels = np.zeros((max_voxels, t, f), dtype=np.float32)
        coors = np.zeros((max_voxels, 3), dtype=np.int32)
        num = np.zeros((max_voxels,), dtype=np.int32)
        m = self.kernel(points, max_voxels, voxels, coors, num)
        return voxels[:m], coors[:m], num[:m]

    def kernel(self, points, max_voxels, voxels, coors, num):
        """Kernel only (variant A of BASELINE.md section 4): grid cached,
        outputs pre-allocated and zeroed by the caller."""
        n, f = points.shape
        gs32 = self._grid_size.astype(np.int32)
        grid = self._dense_grid()
        return _lib().orc_points_to_voxel(
            _p(points, _f32p), n, f, _p(self._voxel_size, _f32p),
            _p(self._point_cloud_range, _f32p), _p(gs32, _i32p),
            self._max_num_points, int(max_voxels), self._overflow,
            _p(grid, _i32p), _p(voxels, _f32p), _p(coors, _i32p), _p(num, _i32p))

    def generate_batched(self, frames, max_voxels=None):
        """Appendix A.5: per-frame voxelization, outputs concatenated in frame
        order, batch index prepended to coords (merge_second_batch contract,
        second/pytorch/models/voxelnet.py:285)."""
        vs, cs, ns, offs = [], [], [], [0]
        for b, pts in enumerate(frames):
            v, c, k = self.generate(pts, max_voxels)
            vs.append(v)
            cs.append(np.concatenate(
                [np.full((c.shape[0], 1), b, dtype=np.int32), c], axis=1))
            ns.append(k)
            offs.append(offs[-1] + v.shape[0])
        return (np.concatenate(vs, 0), np.concatenate(cs, 0), np.concatenate(ns, 0),
                np.asarray(offs, dtype=np.int32))

    @property
    def voxel_size(self):
        return self._voxel_size

    @property
    def point_cloud_range(self):
        return self._point_cloud_range

    @property
    def grid_size(self):
        return self._grid_size

    @property
    def max_num_points_per_voxel(self):
        return self._max_num_points


def merge_sweeps(sweeps):
    """Appendix A.4.  ``sweeps`` = list of (points[N_s,F_in], R[3,3] or None,
    t[3] or None, dt).  Returns the virtual merged cloud [sum N_s, F_in+1]."""
    outs = []
    for pts, R, t, dt in sweeps:
        pts = np.ascontiguousarray(pts, dtype=np.float32)
        n, f_in = pts.shape
        out = np.empty((n, f_in + 1), dtype=np.float32)
        has = R is not None
        Rm = np.ascontiguousarray(R if has else np.eye(3), dtype=np.float32)
        tv = np.ascontiguousarray(t if (has and t is not None) else np.zeros(3), dtype=np.float32)
        _lib().orc_transform_sweep(_p(pts, _f32p), n, f_in, _p(Rm, _f32p), _p(tv, _f32p),
                                   ctypes.c_float(float(np.float32(dt))), int(has),
                                   _p(out, _f32p))
        outs.append(out)
    return np.concatenate(outs, axis=0) if outs else np.zeros((0, 1), np.float32)


def mean_vfe(voxels, num_points, c_out=None):
    voxels = np.ascontiguousarray(voxels, dtype=np.float32)
    num_points = np.ascontiguousarray(num_points, dtype=np.int32)
    m, t, f = voxels.shape
    c_out = min(f, 4) if c_out is None else int(c_out)
    out = np.empty((m, c_out), dtype=np.float32)
    _lib().orc_mean_vfe(_p(voxels, _f32p), _p(num_points, _i32p), m, t, f, c_out, _p(out, _f32p))
    return out


def scatter_bev(feats, coors4, batch_size, d, ny, nx):
    feats = np.ascontiguousarray(feats, dtype=np.float32)
    coors4 = np.ascontiguousarray(coors4, dtype=np.int32)
    m, c = feats.shape
    canvas = np.empty((batch_size, c * d, ny, nx), dtype=np.float32)
    _lib().orc_scatter_bev(_p(feats, _f32p), _p(coors4, _i32p), m, c, batch_size, d, ny, nx,
                           _p(canvas, _f32p))
    return canvas


def sparse_sum_for_anchors_mask(coors, shape):
    """second/core/box_np_ops.py:776-781; ``coors`` [M,3] (z,y,x) or [M,4] (b,z,y,x)."""
    coors = np.ascontiguousarray(coors, dtype=np.int32)
    ny, nx = int(shape[0]), int(shape[1])
    ret = np.empty((ny, nx), dtype=np.float32)
    stride = coors.shape[1]
    _lib().orc_sparse_sum_for_anchors_mask(_p(coors, _i32p), coors.shape[0], stride,
                                           stride - 2, ny, nx, _p(ret, _f32p))
    return ret


def fused_get_anchors_area(dense_map, anchors_bv, stride, offset, grid_sz):
    """second/core/box_np_ops.py:784-810."""
    dense_map = np.ascontiguousarray(dense_map, dtype=np.float32)
    anchors_bv = np.ascontiguousarray(anchors_bv, dtype=np.float32)
    stride = np.ascontiguousarray(stride, dtype=np.float32)
    offset = np.ascontiguousarray(offset, dtype=np.float32)
    grid_sz = np.ascontiguousarray(grid_sz, dtype=np.int64)
    ret = np.empty((anchors_bv.shape[0],), dtype=np.float32)
    _lib().orc_fused_get_anchors_area(_p(dense_map, _f32p), dense_map.shape[0], dense_map.shape[1],
                                      _p(anchors_bv, _f32p), anchors_bv.shape[0],
                                      _p(stride, _f32p), _p(offset, _f32p), _p(grid_sz, _i64p),
                                      _p(ret, _f32p))
    return ret


def anchors_area(coors, anchors_bv, voxel_size, point_cloud_range, grid_sz):
    """The usage pattern of second/kittiviewer/viewer.py:395-401."""
    dm = sparse_sum_for_anchors_mask(coors, tuple(np.asarray(grid_sz)[::-1][1:]))
    _lib().orc_cumsum2d(_p(dm, _f32p), dm.shape[0], dm.shape[1])
    return fused_get_anchors_area(dm, anchors_bv, voxel_size, point_cloud_range, grid_sz)


def crop_mask(points, normal_vec, d):
    """second/core/geometry.py:122-141 (_points_in_convex_polygon_3d_jit for one polygon): float64,
    sign = x*a + y*b + z*c + d summed left to right; a point is inside iff sign < 0 for every
    surface.  ``normal_vec`` [K,3], ``d`` [K] as returned by the reference's surface_equ_3d."""
    p = np.asarray(points, dtype=np.float32)[:, :3].astype(np.float64)
    nv = np.asarray(normal_vec, dtype=np.float64).reshape(-1, 3)
    dd = np.asarray(d, dtype=np.float64).reshape(-1)
    inside = np.ones(p.shape[0], dtype=bool)
    for k in range(nv.shape[0]):
        sign = ((p[:, 0] * nv[k, 0] + p[:, 1] * nv[k, 1]) + p[:, 2] * nv[k, 2]) + dd[k]
        inside &= ~(sign >= 0)
    return inside
