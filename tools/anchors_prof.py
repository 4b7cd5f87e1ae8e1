"""Device time of the anchors-area call on cfg4-shaped output (64 frames x 20000 voxels,
1600 x 1408 map, 70400 anchors) and per kernel, with CUDA events."""
import importlib, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("3dobject_detection_temporal_b200")
B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
c = pkg.synth.CONFIGS["cfg4"]
frames = [pkg.synth.lidar_frame(c["n"], "kitti", i) for i in range(B)]
offs = np.concatenate([[0], np.cumsum([f.shape[0] for f in frames])]).astype(np.int32)
pts = torch.from_numpy(np.concatenate(frames)).cuda()
gen = pkg.VoxelGenerator(c["voxel_size"], c["point_cloud_range"], c["max_num_points"], c["max_voxels"])
_, coors4, _, voff = gen.generate_batched(pts, offs)
rg = c["point_cloud_range"]
xs = np.linspace(rg[0], rg[3], 176, dtype=np.float32); ys = np.linspace(rg[1], rg[4], 200, dtype=np.float32)
cx, cy = np.meshgrid(xs, ys)
bv = np.stack([np.stack([cx - dx / 2, cy - dy / 2, cx + dx / 2, cy + dy / 2], -1).reshape(-1, 4)
               for dx, dy in ((3.9, 1.6), (1.6, 3.9))], 1).reshape(-1, 4).astype(np.float32)
bvd = torch.from_numpy(bv).cuda()
args = (coors4, bvd, gen.voxel_size, gen.point_cloud_range, gen.grid_size)
kw = dict(batch_size=B, max_voxels_per_frame=c["max_voxels"])
for _ in range(3):
    out = pkg.anchors_area(*args, **kw)
torch.cuda.synchronize()
steps = 20
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(steps):
    out = pkg.anchors_area(*args, **kw)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / steps
m = int(coors4.shape[0])
table = B * 1600 * 1408 * 2                       # one write of the 16-bit banded table
alg = table + m * 16 * 2 + out.numel() * 4 + bv.nbytes  # + coors read by both sort passes + areas out + anchors
print(f"frames={B} voxels={m} anchors={bv.shape[0]} call={ms*1e3:.1f}us  algorithmic {alg/1e6:.0f} MB -> "
      f"{alg/ms/1e6:.0f} GB/s  mask true {int((out > 1).sum())}/{out.numel()}")
