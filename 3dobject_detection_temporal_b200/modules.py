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

The three above are parameter-free, so they add no state-dict keys
(torchplus/train/checkpoint.py:129-138 only needs ``.name``).

* ``VFELayer`` / ``VoxelFeatureExtractor`` / ``VoxelFeatureExtractorV2`` and
  ``get_paddings_indicator`` -- voxel_encoder.py:12-189, same constructor arguments and
  the same parameter names (``vfe1.linear.weight``, ``vfe_layers.0.norm.running_mean`` ...),
  so a reference checkpoint loads.  Their bandwidth part runs on the kernels (augmented
  features + padding mask before the first layer, masked max-pool over T at the end:
  ``bvx_vfe_augment`` / ``bvx_masked_max``); the small linear + batch-norm layers in between
  are plain torch modules (the only contraction near this path, C <= 128: library GEMMs).
* ``scatter_index_bev`` -- the camera -> BEV projection scatter of the fusion RPN
  (second/pytorch/models/rpn.py:957-962) for a whole batch in one call.
"""
import torch
from torch import nn

from . import _lib


def _stream(t):
    return torch.cuda.current_stream(t.device).cuda_stream


def _need_cuda(t, what):
    if not t.is_cuda:
        raise RuntimeError(f"{what} must be a CUDA tensor: this module has no CPU fallback")


class _MeanVfeFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, feats, num, c_out, simple_voxel):
        m, t, f = feats.shape
        out = torch.empty((m, 3 if simple_voxel else c_out), dtype=torch.float32, device=feats.device)
        with torch.cuda.device(feats.device):
            _lib.check(_lib.load().bvx_mean_vfe(feats.data_ptr(), num.data_ptr(), m, t, f, c_out,
                                                1 if simple_voxel else 0, out.data_ptr(), _stream(feats)))
        ctx.save_for_backward(feats, num)
        ctx.cfg = (c_out, simple_voxel)
        return out

    @staticmethod
    def backward(ctx, grad):
        # d mean / d features = grad / num, broadcast over T (only reached when something learned
        # precedes the VFE: off the named hot path, so plain torch ops)
        feats, num = ctx.saved_tensors
        c_out, simple_voxel = ctx.cfg
        m, t, f = feats.shape
        nf = num.to(grad.dtype).view(-1, 1)
        g = torch.zeros((m, f), dtype=grad.dtype, device=grad.device)
        if simple_voxel:
            mean = feats[:, :, :4].sum(dim=1) / nf
            r = torch.sqrt(mean[:, 0] ** 2 + mean[:, 1] ** 2).clamp_min(1e-30)
            g[:, 0] = grad[:, 0] * mean[:, 0] / r
            g[:, 1] = grad[:, 0] * mean[:, 1] / r
            g[:, 2] = grad[:, 1]
            g[:, 3] = grad[:, 2]
        else:
            g[:, :c_out] = grad
        return (g / nf).unsqueeze(1).expand(m, t, f).contiguous(), None, None, None


def mean_vfe(features, num_voxels, c_out=None, simple_voxel=False):
    """out[v, c] = sum_t features[v, t, c] / num_voxels[v], c < c_out (default
    min(F, 4), the reference's hard-coded slice, voxel_encoder.py:204).  Computes in fp32 and returns
    the input's dtype; differentiable with respect to ``features``."""
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
    out = _MeanVfeFn.apply(feats, num, c_out, bool(simple_voxel))
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
        # gather of the scattered cells; only needed when a learned VFE feeds the scatter (outside the
        # named hot path), so plain indexing is enough here.  Rows the forward skipped (coordinates
        # outside the canvas) get a zero gradient, like the forward gave them no cell.
        (coors,) = ctx.saved_tensors
        c, d = ctx.dims
        b, _, ny, nx = grad.shape
        g = grad.reshape(b, c, d, ny, nx)
        q = coors.long()
        lim = torch.tensor([b, d, ny, nx], device=q.device)
        ok = ((q >= 0) & (q < lim)).all(dim=1)
        qc = torch.where(ok.unsqueeze(1), q, torch.zeros_like(q))
        out = g[qc[:, 0], :, qc[:, 1], qc[:, 2], qc[:, 3]] * ok.unsqueeze(1).to(grad.dtype)
        return out.contiguous(), None, None, None, None, None


def scatter_bev(voxel_features, coors, batch_size, d, ny, nx):
    """canvas[b, c*D + z, y, x] = voxel_features[i, c] for coors[i] = (b, z, y, x).  Computes in fp32 and
    returns the input's dtype (the reference's scatter_nd keeps it, torchplus/ops/array_ops.py:9).  Rows
    whose coordinates lie outside the canvas are skipped, in the forward and in the backward (the
    reference's index assignment raises on too-large and wraps negative coordinates; the voxelizer never
    produces either)."""
    _need_cuda(voxel_features, "voxel_features")
    if coors.dim() != 2 or coors.shape[1] != 4:
        raise ValueError("coors must be [M, 4] (batch, z, y, x)")
    feats = voxel_features.contiguous().float()
    q = coors.contiguous()
    if q.dtype != torch.int32:
        q = q.int()   # the reference does coors.int() too (middle.py:515)
    canvas = _ScatterFn.apply(feats, q, int(batch_size), int(d), int(ny), int(nx))
    return canvas if voxel_features.dtype == torch.float32 else canvas.to(voxel_features.dtype)


def scatter_index_bev(src, dst_xy, ny, nx, src_idx=None):
    """The camera -> BEV projection scatter of the fusion RPN (second/pytorch/models/rpn.py:957-962) for the
    whole batch in one call: ``out[b, :, dst_xy[b, j, 1], dst_xy[b, j, 0]] = src[b, :, i0[j], i1[j]]`` for
    every j, zeros elsewhere (last j wins where pixels share a cell).  ``src`` [B, C, S0, S1] float32 CUDA;
    ``dst_xy`` [B, P, 2] (x, y) integer; ``src_idx`` [P, 2] or None for the mesh j -> (j // S1, j % S1).
    Forward only."""
    _need_cuda(src, "src")
    if src.dim() != 4 or dst_xy.dim() != 3 or dst_xy.shape[2] != 2 or dst_xy.shape[0] != src.shape[0]:
        raise ValueError("src must be [B, C, S0, S1] and dst_xy [B, P, 2]")
    b, c, s0, s1 = src.shape
    npts = int(dst_xy.shape[1])
    s = src.contiguous().float()
    dxy = dst_xy.to(device=src.device, dtype=torch.int32).contiguous()
    sidx = None
    if src_idx is not None:
        sidx = torch.as_tensor(src_idx).to(device=src.device, dtype=torch.int32).contiguous()
        if tuple(sidx.shape) != (npts, 2):
            raise ValueError("src_idx must be [P, 2]")
    out = torch.empty((b, c, int(ny), int(nx)), dtype=torch.float32, device=src.device)
    with torch.cuda.device(src.device):
        _lib.check(_lib.load().bvx_scatter_index_bev(
            s.data_ptr(), b, c, s0, s1, sidx.data_ptr() if sidx is not None else None, dxy.data_ptr(), npts,
            int(ny), int(nx), out.data_ptr(), _stream(src)))
    return out if src.dtype == torch.float32 else out.to(src.dtype)


# ----------------------------------------------------------------- learned VFEs (voxel_encoder.py:12-189)
def get_paddings_indicator(actual_num, max_num, axis=0):
    """Boolean mask of the real (not padded) slots: ``actual_num[..., None] > arange(max_num)`` along
    ``axis + 1`` (voxel_encoder.py:12-33)."""
    actual_num = torch.unsqueeze(actual_num, axis + 1)
    shape = [1] * actual_num.dim()
    shape[axis + 1] = -1
    steps = torch.arange(max_num, dtype=torch.int, device=actual_num.device).view(shape)
    return actual_num.int() > steps


class _AugmentFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, feats, num, with_distance):
        m, t, f = feats.shape
        fo = f + 3 + (1 if with_distance else 0)
        out = torch.empty((m, t, fo), dtype=torch.float32, device=feats.device)
        mask = torch.empty((m, t), dtype=torch.float32, device=feats.device)
        with torch.cuda.device(feats.device):
            _lib.check(_lib.load().bvx_vfe_augment(feats.data_ptr(), num.data_ptr(), m, t, f,
                                                   1 if with_distance else 0, out.data_ptr(), mask.data_ptr(),
                                                   _stream(feats)))
        ctx.save_for_backward(feats, num)
        ctx.with_distance = with_distance
        ctx.mark_non_differentiable(mask)
        return out, mask

    @staticmethod
    def backward(ctx, g, _gmask):
        feats, num = ctx.saved_tensors
        m, t, f = feats.shape
        gf = g[:, :, :f].clone()
        grel = g[:, :, f:f + 3]
        gf[:, :, :3] += grel - grel.sum(dim=1, keepdim=True) / num.to(g.dtype).view(-1, 1, 1)
        if ctx.with_distance:
            xyz = feats[:, :, :3]
            gf[:, :, :3] += g[:, :, f + 3:f + 4] * xyz / torch.norm(xyz, 2, 2, keepdim=True).clamp_min(1e-30)
        return gf, None, None


class _MaskedMaxFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, num):
        m, t, c = x.shape
        out = torch.empty((m, c), dtype=torch.float32, device=x.device)
        arg = torch.empty((m, c), dtype=torch.int32, device=x.device)
        with torch.cuda.device(x.device):
            _lib.check(_lib.load().bvx_masked_max(x.data_ptr(), num.data_ptr(), m, t, c, out.data_ptr(),
                                                  arg.data_ptr(), _stream(x)))
        ctx.save_for_backward(arg, num)
        ctx.t = t
        return out

    @staticmethod
    def backward(ctx, g):
        arg, num = ctx.saved_tensors
        m, c = g.shape
        a = arg.long()
        live = (a < num.long().view(-1, 1)).to(g.dtype)          # a masked slot passes no gradient (x * 0)
        gx = torch.zeros((m, ctx.t, c), dtype=g.dtype, device=g.device)
        gx.scatter_(1, a.unsqueeze(1), (g * live).unsqueeze(1))
        return gx, None


def vfe_augment(features, num_voxels, with_distance=False):
    """``(cat([features, xyz - points_mean (, |xyz|)], -1), mask[M, T])``: everything VoxelFeatureExtractor / V2
    compute before their first layer (voxel_encoder.py:100-112), in one kernel."""
    _need_cuda(features, "features")
    return _AugmentFn.apply(features.contiguous().float(), num_voxels.contiguous().to(torch.int32), bool(with_distance))


def masked_max(x, num_voxels):
    """``(x * mask).max(dim=1)[0]`` with the padding mask of ``num_voxels`` (voxel_encoder.py:124-129)."""
    _need_cuda(x, "x")
    return _MaskedMaxFn.apply(x.contiguous().float(), num_voxels.contiguous().to(torch.int32))


def _norm_linear(use_norm):
    if use_norm:
        return (lambda c: nn.BatchNorm1d(c, eps=1e-3, momentum=0.01)), (lambda i, o: nn.Linear(i, o, bias=False))
    return (lambda c: nn.Identity()), (lambda i, o: nn.Linear(i, o, bias=True))


def _norm_over_channels(norm, x):
    # BatchNorm1d over the channel axis of [K, T, C] (voxel_encoder.py:54-55)
    return norm(x.permute(0, 2, 1).contiguous()).permute(0, 2, 1).contiguous()


class VFELayer(nn.Module):
    """voxel_encoder.py:36-66: linear -> norm -> relu, concatenated with its max over T."""

    def __init__(self, in_channels, out_channels, use_norm=True, name='vfe'):
        super().__init__()
        self.name = name
        self.units = int(out_channels / 2)
        norm, linear = _norm_linear(use_norm)
        self.linear = linear(in_channels, self.units)
        self.norm = norm(self.units)

    def forward(self, inputs):
        voxel_count = inputs.shape[1]
        pointwise = torch.relu(_norm_over_channels(self.norm, self.linear(inputs)))
        aggregated = torch.max(pointwise, dim=1, keepdim=True)[0]
        return torch.cat([pointwise, aggregated.repeat(1, voxel_count, 1)], dim=2)


class VoxelFeatureExtractor(nn.Module):
    """voxel_encoder.py:68-130.  Same parameters as the reference class; the augmentation + mask before the
    first layer and the masked max over T at the end run on the bvx kernels."""

    def __init__(self, num_input_features=4, use_norm=True, num_filters=(32, 128), with_distance=False,
                 name='VoxelFeatureExtractor'):
        super().__init__()
        self.name = name
        norm, linear = _norm_linear(use_norm)
        assert len(num_filters) == 2
        num_input_features += 3
        if with_distance:
            num_input_features += 1
        self._with_distance = with_distance
        self.vfe1 = VFELayer(num_input_features, num_filters[0], use_norm)
        self.vfe2 = VFELayer(num_filters[0], num_filters[1], use_norm)
        self.linear = linear(num_filters[1], num_filters[1])
        self.norm = norm(num_filters[1])

    def forward(self, features, num_voxels):
        x, mask = vfe_augment(features, num_voxels, self._with_distance)
        mask = mask.unsqueeze(-1)
        x = self.vfe1(x) * mask
        x = self.vfe2(x) * mask
        x = torch.relu(_norm_over_channels(self.norm, self.linear(x)))
        return masked_max(x, num_voxels)


class VoxelFeatureExtractorV2(nn.Module):
    """voxel_encoder.py:132-189: any number of VFE layers."""

    def __init__(self, num_input_features=4, use_norm=True, num_filters=(32, 128), with_distance=False,
                 name='VoxelFeatureExtractor'):
        super().__init__()
        self.name = name
        norm, linear = _norm_linear(use_norm)
        assert len(num_filters) > 0
        num_input_features += 3
        if with_distance:
            num_input_features += 1
        self._with_distance = with_distance
        filters = [num_input_features] + list(num_filters)
        self.vfe_layers = nn.ModuleList([VFELayer(i, o, use_norm) for i, o in zip(filters[:-1], filters[1:])])
        self.linear = linear(filters[-1], filters[-1])
        self.norm = norm(filters[-1])

    def forward(self, features, num_voxels):
        x, mask = vfe_augment(features, num_voxels, self._with_distance)
        mask = mask.unsqueeze(-1)
        for vfe in self.vfe_layers:
            x = vfe(x) * mask
        x = torch.relu(_norm_over_channels(self.norm, self.linear(x)))
        return masked_max(x, num_voxels)


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
    "VoxelFeatureExtractor": VoxelFeatureExtractor,
    "VoxelFeatureExtractorV2": VoxelFeatureExtractorV2,
    "VoxelFeatureExtractorV3": VoxelFeatureExtractorV3,
    "SimpleVoxel": SimpleVoxel,
}
MIDDLE_CLASS_DICT = {
    "PointPillarsScatter": PointPillarsScatter,
}
