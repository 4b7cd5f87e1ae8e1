"""Generate tests/golden/vfe_learned_reference.npz and index_scatter_reference.npz by IMPORTING THE REFERENCE
(this container only; the fixtures are committed).  Run from the repo root:
    python tests/golden/make_golden_vfe.py

Pinned here:
* VoxelFeatureExtractor, VoxelFeatureExtractorV2, get_paddings_indicator -- the reference's own classes
  (second/pytorch/models/voxel_encoder.py:12-189), random parameters and batch-norm statistics, eval mode,
  CPU fp32.  Stored: inputs, every state-dict tensor, the outputs, and the two intermediate tensors the
  kernels produce (augmented features before the first layer, padding mask).
* the camera -> BEV projection assignment of second/pytorch/models/rpn.py:957-962, evaluated on the CPU with the
  reference's own indexing expression (rpn.py itself needs spconv and a GPU to import).
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from make_golden import import_reference  # noqa: E402


def main():
    import torch
    voxel_encoder, _ = import_reference()
    rng = np.random.default_rng(20261105)
    torch.manual_seed(20261105)
    out = {}
    cases = {
        "v1_norm": ("VoxelFeatureExtractor", dict(num_input_features=4, use_norm=True, num_filters=[32, 128], with_distance=False), (301, 35, 4)),
        "v1_dist": ("VoxelFeatureExtractor", dict(num_input_features=4, use_norm=False, num_filters=[16, 64], with_distance=True), (97, 5, 4)),
        "v2_three": ("VoxelFeatureExtractorV2", dict(num_input_features=5, use_norm=True, num_filters=[16, 32, 64], with_distance=True), (130, 10, 5)),
    }
    for tag, (cls, kw, (m, t, f)) in cases.items():
        net = getattr(voxel_encoder, cls)(**kw)
        # random batch-norm statistics, then eval mode
        for mod in net.modules():
            if isinstance(mod, torch.nn.BatchNorm1d):
                mod.running_mean.normal_(0, 0.5)
                mod.running_var.uniform_(0.5, 2.0)
                mod.weight.data.uniform_(0.5, 1.5)
                mod.bias.data.normal_(0, 0.2)
        net.eval()
        num = rng.integers(1, t + 1, size=m).astype(np.int32)
        vox = rng.normal(0, 10, size=(m, t, f)).astype(np.float32)
        for v in range(m):
            vox[v, num[v]:] = 0
        feats, nv = torch.from_numpy(vox), torch.from_numpy(num)
        with torch.no_grad():
            y = net(feats, nv).numpy()
            # the intermediate tensors, with the reference's own expressions (voxel_encoder.py:100-112)
            mean = feats[:, :, :3].sum(dim=1, keepdim=True) / nv.type_as(feats).view(-1, 1, 1)
            rel = feats[:, :, :3] - mean
            parts = [feats, rel] + ([torch.norm(feats[:, :, :3], 2, 2, keepdim=True)] if kw["with_distance"] else [])
            aug = torch.cat(parts, dim=-1).numpy()
            mask = voxel_encoder.get_paddings_indicator(nv, t, axis=0).numpy()
        out[f"{tag}_voxels"], out[f"{tag}_num"], out[f"{tag}_out"] = vox, num, y
        out[f"{tag}_aug"], out[f"{tag}_mask"] = aug, mask
        for k, v in net.state_dict().items():
            out[f"{tag}_sd_{k}"] = v.numpy()
    np.savez_compressed(os.path.join(HERE, "vfe_learned_reference.npz"), **out)

    # --- camera -> BEV projection (rpn.py:949-962): mesh of (c, r), per-batch target indices, last write wins
    out = {}
    for tag, (b, c, s0, s1, ny, nx) in {"fusion": (2, 6, 26, 8, 50, 44), "dup": (1, 3, 12, 4, 8, 8)}.items():
        cc, rr = np.meshgrid(np.arange(s0), np.arange(s1))        # n_mesh(np.arange(156), np.arange(48)) in the reference
        points = np.stack([cc, rr]).reshape((2, -1)).T
        src = rng.normal(0, 1, size=(b, c, s0, s1)).astype(np.float32)
        idxs = np.stack([rng.integers(0, nx, size=(b, points.shape[0])),
                         rng.integers(0, ny, size=(b, points.shape[0]))], axis=-1).astype(np.int32)
        projected = torch.zeros(b, c, ny, nx)
        fuse = torch.from_numpy(src)
        for b_idx in range(b):
            i = torch.from_numpy(idxs[b_idx])
            projected[b_idx, :, i[:, 1].to(torch.int64), i[:, 0].to(torch.int64)] = \
                fuse[b_idx][:, points[:, 0].astype(np.int64), points[:, 1].astype(np.int64)]
        out[f"{tag}_src"], out[f"{tag}_points"], out[f"{tag}_idxs"] = src, points.astype(np.int32), idxs
        out[f"{tag}_out"] = projected.numpy()
    np.savez_compressed(os.path.join(HERE, "index_scatter_reference.npz"), **out)
    print("wrote vfe_learned_reference.npz, index_scatter_reference.npz")


if __name__ == "__main__":
    main()
