"""Short driver for ncu captures: N steps of generate_batched on a cfg4-shaped batch.

usage: python tools/ncu_step.py [frames=64] [overflow=break] [steps=8] [cfg=cfg4]
"""
import importlib
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("3dobject_detection_temporal_b200")
frames_n = int(sys.argv[1]) if len(sys.argv) > 1 else 64
overflow = sys.argv[2] if len(sys.argv) > 2 else "break"
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 8
cfg = sys.argv[4] if len(sys.argv) > 4 else "cfg4"
c = pkg.synth.CONFIGS[cfg]
frames = [pkg.synth.lidar_frame(c["n"], c["sensor"], c["seed0"] + i) for i in range(frames_n)]
offs = np.concatenate([[0], np.cumsum([f.shape[0] for f in frames])]).astype(np.int32)
sets = [torch.from_numpy(np.concatenate([frames[(i + 7 * k) % frames_n] for i in range(frames_n)])).cuda()
        for k in range(3)]
gen = pkg.VoxelGenerator(c["voxel_size"], c["point_cloud_range"], c["max_num_points"], c["max_voxels"],
                         overflow=overflow)
out = gen.generate_batched(sets[0], offs, mean_channels=4, padded=True)
for i in range(steps):
    gen.generate_batched(sets[i % 3], offs, mean_channels=4, padded=True, out=out)
torch.cuda.synchronize()
print("ok", int(out[3][-1].item()))
