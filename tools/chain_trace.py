"""Phase breakdown of k_chain from its per-ticket globaltimer stamps.

usage: python tools/chain_trace.py [frames=64] [overflow=break] [cfg=cfg4]
"""
import importlib
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("3dobject_detection_temporal_b200")
B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
overflow = sys.argv[2] if len(sys.argv) > 2 else "break"
cfg = sys.argv[3] if len(sys.argv) > 3 else "cfg4"
c = pkg.synth.CONFIGS[cfg]
frames = [pkg.synth.lidar_frame(c["n"], c["sensor"], c["seed0"] + i) for i in range(B)]
offs = np.concatenate([[0], np.cumsum([f.shape[0] for f in frames])]).astype(np.int32)
pts = torch.from_numpy(np.concatenate(frames)).cuda()
gen = pkg.VoxelGenerator(c["voxel_size"], c["point_cloud_range"], c["max_num_points"], c["max_voxels"],
                         overflow=overflow)
out = gen.generate_batched(pts, offs, mean_channels=4, padded=True)
gen.chain_trace(1 << 16)
for _ in range(3):
    gen.generate_batched(pts, offs, mean_channels=4, padded=True, out=out)
torch.cuda.synchronize()
tr, cp = gen.chain_trace()
rows = tr.shape[0] // B
tr = tr.reshape(rows, B, 8).astype(np.float64)
t0 = tr[tr > 0].min()
names = ["prep(0-1)", "wait(1-2)", "resolve(2-3)", "number(3-4)", "publish(4-5)", "bucket(5-6)"]
print(f"B={B} {overflow} {cfg}: chunk={cp} points, rows={rows}; kernel span {(tr.max() - t0) / 1e3:.1f} us")
for f in (0, B // 2, B - 1):
    print(f" frame {f}:")
    for r in range(rows):
        s = tr[r, f]
        if s[0] == 0:
            continue
        if s[5] == 0:
            print(f"  chunk {r:3d}: taken at {(s[0]-t0)/1e3:7.1f} us, skipped (closed)")
            continue
        d = np.diff(s[:7]) / 1e3
        print(f"  chunk {r:3d}: taken {(s[0]-t0)/1e3:7.1f}  " + "  ".join(f"{n} {v:6.2f}" for n, v in zip(names, d))
              + f"   handed at {(s[5]-t0)/1e3:7.1f}")
ok = tr[:, :, 5] > 0
d = np.diff(tr[:, :, :7], axis=2) / 1e3
print(" mean over processed chunks:", "  ".join(f"{n} {d[:, :, i][ok].mean():6.2f}" for i, n in enumerate(names)))
step = []
for f in range(B):
    h = tr[:, f, 5]
    h = h[h > 0]
    if len(h) > 1:
        step.append(np.diff(h).mean() / 1e3)
print(f" mean hand-over to hand-over interval per frame: {np.mean(step):.2f} us")
