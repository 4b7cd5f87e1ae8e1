"""Generate tests/golden/*.npz by IMPORTING THE REFERENCE (this container only).

Run from the repo root:  python tests/golden/make_golden.py
Needs /root/reference (read-only).  The fixtures it writes are committed, so
nothing at test/bench/smoke time reads /root/reference.

What is pinned here:
* VoxelFeatureExtractorV3.forward and SimpleVoxel.forward
  (second/pytorch/models/voxel_encoder.py:191-228), imported under two shims
  applied before import: a stub ``spconv`` module (imported but unused by these
  classes, voxel_encoder.py:6) and ``collections.Iterable`` (py3.12 removed it;
  torchplus/train/optim.py:1).
* torchplus.scatter_nd (torchplus/ops/array_ops.py:7-20) used the way a BEV
  scatter uses it: indices (b, y, x) into a [B, H, W, C] canvas, then permuted to
  NCHW -- and the SparseConvTensor.dense()-style [N, C*D, H, W] view of
  second/pytorch/models/middle.py:522-525 for D > 1.
The voxelizer itself cannot be pinned this way: its source is not in the tree.
"""
import collections
import collections.abc
import os
import sys
import types
import warnings

import numpy as np

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))


def import_reference():
    sys.modules.setdefault("spconv", types.ModuleType("spconv"))
    collections.Iterable = collections.abc.Iterable
    sys.path.insert(0, REF)
    from second.pytorch.models import voxel_encoder  # noqa
    import torchplus  # noqa
    return voxel_encoder, torchplus


def main():
    import torch
    voxel_encoder, torchplus = import_reference()
    rng = np.random.default_rng(20261104)

    # --- mean VFE golden: ragged counts, zero padding, T=5/F=4 and T=100/F=4, F=5 (drops dt)
    out = {}
    for tag, (m, t, f) in {"t5f4": (257, 5, 4), "t100f4": (33, 100, 4), "t10f5": (64, 10, 5)}.items():
        num = rng.integers(1, t + 1, size=m).astype(np.int32)
        vox = (rng.normal(0, 20, size=(m, t, f))).astype(np.float32)
        for v in range(m):
            vox[v, num[v]:] = 0
        vfe = voxel_encoder.VoxelFeatureExtractorV3(num_input_features=f)
        sv = voxel_encoder.SimpleVoxel(num_input_features=f)
        with torch.no_grad():
            y = vfe(torch.from_numpy(vox), torch.from_numpy(num)).numpy()
            ys = sv(torch.from_numpy(vox), torch.from_numpy(num)).numpy()
        out[f"{tag}_voxels"] = vox
        out[f"{tag}_num"] = num
        out[f"{tag}_v3"] = y
        out[f"{tag}_simple"] = ys
    np.savez_compressed(os.path.join(HERE, "vfe_reference.npz"), **out)

    # --- scatter golden
    out = {}
    for tag, (b, c, d, h, w, m) in {"pp_c4": (2, 4, 1, 31, 29, 300),
                                     "pp_c64": (1, 64, 1, 16, 24, 100),
                                     "d2_c3": (2, 3, 2, 9, 11, 120)}.items():
        cells = rng.choice(b * d * h * w, size=m, replace=False)
        bb, rem = np.divmod(cells, d * h * w)
        zz, rem = np.divmod(rem, h * w)
        yy, xx = np.divmod(rem, w)
        coors = np.stack([bb, zz, yy, xx], axis=1).astype(np.int32)
        feats = rng.normal(0, 1, size=(m, c)).astype(np.float32)
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            dense = torchplus.scatter_nd(torch.from_numpy(coors).long(), torch.from_numpy(feats),
                                         [b, d, h, w, c])
        # SparseConvTensor.dense(): [N, D, H, W, C] -> [N, C, D, H, W]; then view [N, C*D, H, W]
        canvas = dense.permute(0, 4, 1, 2, 3).contiguous().view(b, c * d, h, w).numpy()
        out[f"{tag}_coors"] = coors
        out[f"{tag}_feats"] = feats
        out[f"{tag}_canvas"] = canvas
        out[f"{tag}_shape"] = np.array([b, c, d, h, w], dtype=np.int64)
    np.savez_compressed(os.path.join(HERE, "scatter_reference.npz"), **out)
    print("wrote", sorted(os.listdir(HERE)))


if __name__ == "__main__":
    main()
