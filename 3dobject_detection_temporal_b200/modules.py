"""nn.Module mirrors of the reference's voxel feature extractor and middle
extractor protocols, running on the kernels behind include/bvx.h.

* ``VoxelFeatureExtractorV3`` / ``SimpleVoxel`` -- second/pytorch/models/
  voxel_encoder.py:191-228: ``forward(features[M,T,F], num_voxels[M]) -> [M,C]``.
  Constructor arguments are accepted and ignored exactly like the reference does
  (voxel_encoder.py:192-199).  Looked up by class-name string in the detector's
  registry (second/pytorch/models/voxelnet.py:129-135).
* ``PointPillarsScatter`` -- satisfies the middle-extractor protocol
  (ctor ``(output_shape, use_norm, num_input_features, num_filters_down1,
  num_filters_down2)``, second/pytorch/models/voxelnet.py:184-189;
  ``forward(voxel_features, coors, batch_size) -> [N, C*D, H, W]``,
  second/pytorch/models/middle.py:513-526) with the math of
  torchplus.scatter_nd (torchplus/ops/array_ops.py:7-20).

All three are parameter-free, so they add no state-dict keys
(torchplus/train/checkpoint.py:129-138 only needs ``.name``).
"""
import torch
from torch import nn

from . import _lib


def _stream(t):
    return torch.cuda.current_stream(t.device).cuda_stream


def _need_cuda(t, what):
    if not t.is_cuda:
        raise RuntimeError(f"{what} must be a CUDA tensor: this module has no CPU fallback")


def mean_vfe(features, num_voxels, c_out=None, simple_voxel=False):
    """out[v, c] = sum_t features[v, t, c] / num_voxels[v], c < c_out (default
    min(F, 4), the reference's hard-coded slice, voxel_encoder.py:204)."""
    _need_cuda(features, "features")
    if features.dim() != 3:
        raise ValueError("features must be [M, T, F]")
    m, t, f = features.shape
    c_out = min(f, 4) if c_out is None else int(c_out)
    feats = features.contiguous()
    if feats.dtype != torch.float32:
        feats = feats.float()
    num = num_voxels.contiguous()
    if num.dtype != torch.int32:
        num = num.to(torch.int32)
    out = torch.empty((m, 3 if simple_voxel else c_out), dtype=torch.float32, device=feats.device)
    with torch.cuda.device(feats.device):
        _lib.check(_lib.load().bvx_mean_vfe(feats.data_ptr(), num.data_ptr(), m, t, f, c_out,
                                            1 if simple_voxel else 0, out.data_ptr(), _stream(feats)))
    return out if features.dtype == torch.float32 else out.to(features.dtype)


class VoxelFeatureExtractorV3(nn.Module):
    """Mean VFE.  ``num_output_channels=None`` keeps the reference's 4-channel
    slice; pass F to keep the dt channel of temporal input (north star)."""

    def __init__(self, num_input_features=4, use_norm=True, num_filters=(32, 128),
                 with_distance=False, name='VoxelFeatureExtractor', num_output_channels=None):
        super().__init__()
        self.name = name
        self.num_output_channels = num_output_channels

    def forward(self, features, num_voxels):
        return mean_vfe(features, num_voxels, self.num_output_channels)


class SimpleVoxel(nn.Module):
    """Mean VFE followed by [ |xy|_2, z, reflectance ] (voxel_encoder.py:208-228)."""

    def __init__(self, num_input_features=4, use_norm=True, num_filters=(32, 128),
                 with_distance=False, name='SimpleVoxel'):
        super().__init__()
        self.name = name

    def forward(self, features, num_voxels):
        return mean_vfe(features, num_voxels, 4, simple_voxel=True)


class _ScatterFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, feats, coors, b, d, ny, nx):
        m, c = feats.shape
        canvas = torch.empty((b, c * d, ny, nx), dtype=torch.float32, device=feats.device)
        with torch.cuda.device(feats.device):
            _lib.check(_lib.load().bvx_scatter_bev(feats.data_ptr(), coors.data_ptr(), m, c, b, d, ny, nx,
                                                   canvas.data_ptr(), _stream(feats)))
        ctx.save_for_backward(coors)
        ctx.dims = (c, d)
        return canvas

    @staticmethod
    def backward(ctx, grad):
        # gather of the scattered cells; only needed when a learned VFE feeds the
        # scatter (outside the named hot path), so plain indexing is enough here
        (coors,) = ctx.saved_tensors
        c, d = ctx.dims
        b, _, ny, nx = grad.shape
        g = grad.view(b, c, d, ny, nx)
        q = coors.long()
        return g[q[:, 0], :, q[:, 1], q[:, 2], q[:, 3]].contiguous(), None, None, None, None, None


def scatter_bev(voxel_features, coors, batch_size, d, ny, nx):
    """canvas[b, c*D + z, y, x] = voxel_features[i, c] for coors[i] = (b, z, y, x)."""
    _need_cuda(voxel_features, "voxel_features")
    if coors.dim() != 2 or coors.shape[1] != 4:
        raise ValueError("coors must be [M, 4] (batch, z, y, x)")
    feats = voxel_features.contiguous().float()
    q = coors.contiguous()
    if q.dtype != torch.int32:
        q = q.int()   # the reference does coors.int() too (middle.py:515)
    return _ScatterFn.apply(feats, q, int(batch_size), int(d), int(ny), int(nx))


class PointPillarsScatter(nn.Module):
    """BEV canvas scatter behind the middle-extractor protocol.

    ``output_shape`` is the detector's ``dense_shape = [1, D, H, W, C]``
    (second/pytorch/builder/second_builder.py:18)."""

    def __init__(self, output_shape, use_norm=True, num_input_features=64,
                 num_filters_down1=(64,), num_filters_down2=(64, 64), name='PointPillarsScatter'):
        super().__init__()
        self.name = name
        self.output_shape = list(output_shape)
        self.nz, self.ny, self.nx = (int(v) for v in output_shape[1:4])
        self.num_input_features = num_input_features

    def forward(self, voxel_features, coors, batch_size):
        return scatter_bev(voxel_features, coors, batch_size, self.nz, self.ny, self.nx)


# class-name registries in the style of second/pytorch/models/voxelnet.py:129-135,171-183
VFE_CLASS_DICT = {
    "VoxelFeatureExtractorV3": VoxelFeatureExtractorV3,
    "SimpleVoxel": SimpleVoxel,
}
MIDDLE_CLASS_DICT = {
    "PointPillarsScatter": PointPillarsScatter,
}
