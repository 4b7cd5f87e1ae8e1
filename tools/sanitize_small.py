"""Small pass over every entry point, meant to be run under `compute-sanitizer --tool memcheck`:
single frame, sweeps with a shuffle permutation, batched device form with the fused mean VFE, the
pipelined host form in both assembly modes (compact rows, point indices) and with device rows, the
fused points -> canvas path, the anchors area."""
import importlib, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("3dobject_detection_temporal_b200")
synth = pkg.synth
K = dict(voxel_size=[0.2, 0.2, 0.4], point_cloud_range=[0, -40, -3, 70.4, 40, 1])
frames = [synth.lidar_frame(n, "kitti", seed=i) for i, n in enumerate([6000, 0, 2500, 1, 4000])]
offs = np.concatenate([[0], np.cumsum([f.shape[0] for f in frames])]).astype(np.int32)
cat = np.concatenate(frames)
for overflow in ("break", "continue"):
    for T in (5, 12):
        g = pkg.VoxelGenerator(K["voxel_size"], K["point_cloud_range"], T, 700, overflow=overflow)
        g.generate(torch.from_numpy(frames[0]).cuda())
        perm = torch.randperm(frames[2].shape[0]).int().cuda()
        g.generate_sweeps([(torch.from_numpy(frames[0]).cuda(), None, None, 0.0),
                           (torch.from_numpy(frames[2]).cuda(), np.eye(3, dtype=np.float32), np.zeros(3, np.float32), 0.1, perm)])
        g.generate_batched(torch.from_numpy(cat).cuda(), offs, mean_channels=4)
        g.generate_bev(torch.from_numpy(cat).cuda(), offs)
        g.close()
        for asm in (True, False):
            gh = pkg.VoxelGenerator(K["voxel_size"], K["point_cloud_range"], T, 700, overflow=overflow, host_assemble=asm)
            hp = pkg.pinned_empty(cat.shape, np.float32); hp[:] = cat
            b = len(frames)
            out = (pkg.pinned_empty((b * 700, T, 4), np.float32), pkg.pinned_empty((b * 700, 4), np.int32),
                   pkg.pinned_empty((b * 700,), np.int32), pkg.pinned_empty((b + 1,), np.int32))
            for _ in range(2):
                gh.generate_batched(hp, offs, out=out)
            gh.close()
torch.cuda.synchronize()
print("sanitize_small: done")
