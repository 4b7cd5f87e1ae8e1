"""ORACLE -- test infrastructure.  numba nopython twin of the reference's absent
second/core/point_cloud/point_cloud_ops.py::_points_to_voxel_reverse_kernel
(signature evidence: second/kittiviewer/viewer.py:37,385-392), restated from
SURVEY.md Appendix A.3.  Optional: only used when numba imports.
"""
import numpy as np

try:
    import numba
    HAVE_NUMBA = True
except Exception:  # pragma: no cover
    numba = None
    HAVE_NUMBA = False

if HAVE_NUMBA:
    @numba.njit(cache=False)
    def _kernel(points, voxel_size, coors_range, grid_size, num_points_per_voxel,
                coor_to_voxelidx, voxels, coors, max_points, max_voxels, do_break):
        n = points.shape[0]
        coor = np.zeros((3,), dtype=np.int32)
        voxel_num = 0
        for i in range(n):
            failed = False
            for j in range(3):
                c = np.floor((points[i, j] - coors_range[j]) / voxel_size[j])
                if not (c >= 0 and c < grid_size[j]):
                    failed = True
                    break
                coor[2 - j] = np.int32(c)
            if failed:
                continue
            voxelidx = coor_to_voxelidx[coor[0], coor[1], coor[2]]
            if voxelidx == -1:
                if voxel_num >= max_voxels:
                    if do_break:
                        break
                    continue
                voxelidx = voxel_num
                voxel_num += 1
                coor_to_voxelidx[coor[0], coor[1], coor[2]] = voxelidx
                coors[voxelidx, 0] = coor[0]
                coors[voxelidx, 1] = coor[1]
                coors[voxelidx, 2] = coor[2]
            num = num_points_per_voxel[voxelidx]
            if num < max_points:
                voxels[voxelidx, num] = points[i]
                num_points_per_voxel[voxelidx] = num + 1
        for v in range(voxel_num):
            coor_to_voxelidx[coors[v, 0], coors[v, 1], coors[v, 2]] = -1
        return voxel_num


def points_to_voxel(points, voxel_size, coors_range, max_points, max_voxels,
                    overflow="break", grid=None):
    if not HAVE_NUMBA:
        raise RuntimeError("numba is not importable")
    points = np.ascontiguousarray(points, dtype=np.float32)
    vs = np.asarray(voxel_size, dtype=np.float32)
    rg = np.asarray(coors_range, dtype=np.float32)
    gs = np.round((rg[3:] - rg[:3]) / vs).astype(np.int64)
    if grid is None:
        grid = np.full(tuple(gs[::-1]), -1, dtype=np.int32)
    n, f = points.shape
    voxels = np.zeros((max_voxels, max_points, f), dtype=np.float32)
    coors = np.zeros((max_voxels, 3), dtype=np.int32)
    num = np.zeros((max_voxels,), dtype=np.int32)
    m = _kernel(points, vs, rg, gs, num, grid, voxels, coors, max_points, max_voxels,
                overflow == "break")
    return voxels[:m], coors[:m], num[:m]
