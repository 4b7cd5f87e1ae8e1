"""Anchors mask from voxel coordinates on the device.

The reference computes, in the CPU worker right after ``generate()``
(pattern of second/kittiviewer/viewer.py:395-401)::

    dense = box_np_ops.sparse_sum_for_anchors_mask(coors, tuple(grid_size[::-1][1:]))
    dense = dense.cumsum(0).cumsum(1)
    area  = box_np_ops.fused_get_anchors_area(dense, anchors_bv, voxel_size, pc_range, grid_size)
    anchors_mask = area > anchor_area_threshold

``anchors_area`` / ``anchors_mask`` do all of it in one C-ABI call for a whole
batch, taking the ``coors`` tensor ``VoxelGenerator.generate[_batched]`` left on
the device.  ``sparse_sum_for_anchors_mask`` and ``fused_get_anchors_area`` are
the two reference functions under their own names and signatures
(second/core/box_np_ops.py:776-781, :784-810) for callers that keep the
decomposition.  CUDA tensors only: there is no CPU fallback.
"""
import ctypes

import numpy as np
import torch

from . import _lib

_WS = {}   # (device index) -> scratch tensor, grown on demand


def _stream(t):
    return torch.cuda.current_stream(t.device).cuda_stream


def _need_cuda(t, what):
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise RuntimeError(f"{what} must be a CUDA tensor: this module has no CPU fallback")


def _f2(a, i0):
    a = np.asarray(a, dtype=np.float32).reshape(-1)
    return (ctypes.c_float * 2)(float(a[i0]), float(a[i0 + 1]))


def _i2(a):
    a = np.asarray(a).reshape(-1)
    return (ctypes.c_int32 * 2)(int(a[0]), int(a[1]))


def _as_i32(coors):
    q = coors.contiguous()
    return q if q.dtype == torch.int32 else q.int()


def _as_f32(a, device):
    if not isinstance(a, torch.Tensor):
        a = torch.as_tensor(np.ascontiguousarray(a, dtype=np.float32))
    return a.to(device=device, dtype=torch.float32).contiguous()


def anchors_area(coors, anchors_bv, voxel_size, point_cloud_range, grid_size, batch_size=None,
                 max_voxels_per_frame=None):
    """Voxel count under every bird's-eye-view anchor box.

    coors: CUDA int32 [M, 3] (z, y, x) or [M, 4] (batch, z, y, x); anchors_bv [A, 4]
    (x0, y0, x1, y1) shared by all frames.  Returns float32 [A] for [M, 3] input
    (the reference's shape) or [batch_size, A] for [M, 4]."""
    _need_cuda(coors, "coors")
    if coors.dim() != 2 or coors.shape[1] not in (3, 4):
        raise ValueError("coors must be [M, 3] (z, y, x) or [M, 4] (batch, z, y, x)")
    q = _as_i32(coors)
    m, cols = q.shape
    if cols == 3:
        if batch_size not in (None, 1):
            raise ValueError("[M, 3] coordinates carry no batch index")
        b = 1
    else:
        if batch_size is None:
            raise ValueError("batch_size is required with [M, 4] coordinates")
        b = int(batch_size)
    anchors = _as_f32(anchors_bv, q.device)
    if anchors.dim() != 2 or anchors.shape[1] != 4:
        raise ValueError("anchors_bv must be [A, 4] (x0, y0, x1, y1)")
    a = anchors.shape[0]
    grid = _i2(grid_size)
    cap = int(m if max_voxels_per_frame is None else max_voxels_per_frame)
    lib = _lib.load()
    need = lib.bvx_anchors_area_workspace_bytes(b, grid[1], grid[0], cap, m)
    if need == 0:
        raise ValueError("bad or unsupported anchors-area shape")
    key = q.device.index
    ws = _WS.get(key)
    if ws is None or ws.numel() < need:
        ws = _WS[key] = torch.empty(need, dtype=torch.uint8, device=q.device)
    out = torch.empty((b, a), dtype=torch.float32, device=q.device)
    with torch.cuda.device(q.device):
        _lib.check(lib.bvx_anchors_area(q.data_ptr(), m, cols, b, _f2(voxel_size, 0), _f2(point_cloud_range, 0),
                                        grid, cap, anchors.data_ptr(), a, ws.data_ptr(), ws.numel(),
                                        out.data_ptr(), _stream(q)))
    return out[0] if cols == 3 else out


def anchors_mask(coors, anchors_bv, voxel_size, point_cloud_range, grid_size, anchor_area_threshold=1,
                 batch_size=None, max_voxels_per_frame=None):
    """``anchors_area(...) > anchor_area_threshold`` (second/kittiviewer/viewer.py:399-401;
    threshold knob second/protos/input_reader.proto:23)."""
    return anchors_area(coors, anchors_bv, voxel_size, point_cloud_range, grid_size, batch_size,
                        max_voxels_per_frame) > anchor_area_threshold


def sparse_sum_for_anchors_mask(coors, shape):
    """second/core/box_np_ops.py:776-781: float32 [ny, nx] histogram of coors[:, -2:]."""
    _need_cuda(coors, "coors")
    if coors.dim() != 2 or coors.shape[1] not in (3, 4):
        raise ValueError("coors must be [M, 3] or [M, 4]")
    q = _as_i32(coors)
    ny, nx = int(shape[0]), int(shape[1])
    ret = torch.empty((ny, nx), dtype=torch.float32, device=q.device)
    with torch.cuda.device(q.device):
        _lib.check(_lib.load().bvx_sparse_sum_for_anchors_mask(q.data_ptr(), q.shape[0], q.shape[1], ny, nx,
                                                               ret.data_ptr(), _stream(q)))
    return ret


def fused_get_anchors_area(dense_map, anchors_bv, stride, offset, grid_size):
    """second/core/box_np_ops.py:784-810 on a summed-area table the caller built."""
    _need_cuda(dense_map, "dense_map")
    if dense_map.dim() != 2:
        raise ValueError("dense_map must be [ny, nx]")
    dm = dense_map.contiguous().float()
    anchors = _as_f32(anchors_bv, dm.device)
    if anchors.dim() != 2 or anchors.shape[1] != 4:
        raise ValueError("anchors_bv must be [A, 4]")
    ret = torch.empty((anchors.shape[0],), dtype=torch.float32, device=dm.device)
    with torch.cuda.device(dm.device):
        _lib.check(_lib.load().bvx_fused_get_anchors_area(dm.data_ptr(), dm.shape[0], dm.shape[1],
                                                          anchors.data_ptr(), anchors.shape[0], _f2(stride, 0),
                                                          _f2(offset, 0), _i2(grid_size), ret.data_ptr(),
                                                          _stream(dm)))
    return ret
