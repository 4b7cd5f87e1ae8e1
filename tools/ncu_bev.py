"""Short driver for ncu captures of the pillar path: cfg3 fused points -> canvas (C = 4) and the BEV scatter alone
with C = 64 features per pillar.  usage: python tools/ncu_bev.py [steps=4]"""
import importlib, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("3dobject_detection_temporal_b200")
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 4
c = pkg.synth.CONFIGS["cfg3"]
pts = torch.from_numpy(pkg.synth.lidar_frame(c["n"], c["sensor"], c["seed0"])).cuda()
offs = np.asarray([0, c["n"]], np.int32)
gen = pkg.VoxelGenerator(c["voxel_size"], c["point_cloud_range"], c["max_num_points"], c["max_voxels"])
out = None
for _ in range(steps):
    out = gen.generate_bev(pts, offs, channels=4, out=out)
canvas, feats, coors, num, fvo = out
torch.cuda.synchronize()
m = int(fvo[-1].item())
gx, gy, gz = (int(x) for x in gen.grid_size)
f64 = torch.randn((m, 64), device="cuda")
sc = pkg.PointPillarsScatter([1, gz, gy, gx, 64])
for _ in range(steps):
    big = sc(f64, coors[:m], 1)
torch.cuda.synchronize()
print("ok", m, tuple(canvas.shape), tuple(big.shape))
