"""ORACLE -- test infrastructure.  Pure-Python/numpy restatement of SURVEY.md
Appendix A (tiny cases only; the C kernel in points_to_voxel.c is the workhorse).

Independent of the C file on purpose: uses a dict instead of the dense grid and
numpy float32 scalars for the cell arithmetic, so a slip in either shows up as a
disagreement in tests/test_oracle.py.
"""
import numpy as np


def grid_size(voxel_size, point_cloud_range):
    vs = np.asarray(voxel_size, dtype=np.float32)
    rg = np.asarray(point_cloud_range, dtype=np.float32)
    return np.round((rg[3:] - rg[:3]) / vs).astype(np.int64)


def points_to_voxel(points, voxel_size, coors_range, max_points, max_voxels,
                    overflow="break"):
    """(voxels, coors, num_points) -- wrapper signature pinned by
    second/kittiviewer/viewer.py:385-392."""
    points = np.asarray(points, dtype=np.float32)
    vs = np.asarray(voxel_size, dtype=np.float32)
    rg = np.asarray(coors_range, dtype=np.float32)
    gs = grid_size(vs, rg)
    n, f = points.shape
    voxels = np.zeros((max_voxels, max_points, f), dtype=np.float32)
    coors = np.zeros((max_voxels, 3), dtype=np.int32)
    num = np.zeros((max_voxels,), dtype=np.int32)
    table = {}
    voxel_num = 0
    for i in range(n):
        coor = [0, 0, 0]
        ok = True
        for j in range(3):
            c = np.floor((points[i, j] - rg[j]) / vs[j])  # float32 throughout
            assert c.dtype == np.float32
            if not (c >= 0 and c < gs[j]):                # NaN -> dropped (defined here)
                ok = False
                break
            coor[2 - j] = int(c)
        if not ok:
            continue
        key = tuple(coor)
        vid = table.get(key, -1)
        if vid == -1:
            if voxel_num >= max_voxels:
                if overflow == "break":
                    break
                continue
            vid = voxel_num
            voxel_num += 1
            table[key] = vid
            coors[vid] = coor
        if num[vid] < max_points:
            voxels[vid, num[vid]] = points[i]
            num[vid] += 1
    return voxels[:voxel_num], coors[:voxel_num], num[:voxel_num]


def merge_sweeps(sweeps):
    """Appendix A.4 in numpy elementwise fp32 (fixed association, no FMA)."""
    outs = []
    for pts, R, t, dt in sweeps:
        pts = np.asarray(pts, dtype=np.float32)
        n, f_in = pts.shape
        out = np.empty((n, f_in + 1), dtype=np.float32)
        if R is None:
            out[:, :3] = pts[:, :3]
        else:
            R = np.asarray(R, dtype=np.float32)
            t = np.asarray(t, dtype=np.float32)
            x, y, z = pts[:, 0], pts[:, 1], pts[:, 2]
            for r in range(3):
                out[:, r] = ((R[r, 0] * x + R[r, 1] * y) + R[r, 2] * z) + t[r]
        out[:, 3:f_in] = pts[:, 3:]
        out[:, f_in] = np.float32(dt)
        outs.append(out)
    return np.concatenate(outs, axis=0)


def mean_vfe(voxels, num_points, c_out=None):
    voxels = np.asarray(voxels, dtype=np.float32)
    m, t, f = voxels.shape
    c_out = min(f, 4) if c_out is None else c_out
    s = np.zeros((m, c_out), dtype=np.float32)
    for k in range(t):
        s = s + voxels[:, k, :c_out]
    return s / np.asarray(num_points).astype(np.float32).reshape(-1, 1)


def scatter_bev(feats, coors4, batch_size, d, ny, nx):
    feats = np.asarray(feats, dtype=np.float32)
    m, c = feats.shape
    canvas = np.zeros((batch_size, c, d, ny, nx), dtype=np.float32)
    for i in range(m):
        b, z, y, x = (int(v) for v in coors4[i])
        canvas[b, :, z, y, x] = feats[i]
    return canvas.reshape(batch_size, c * d, ny, nx)
