"""Per-kernel device times inside the pinned batched host form (sub-batches), plus wall time per call."""
import importlib, os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("3dobject_detection_temporal_b200")
B = 64
c = pkg.synth.CONFIGS["cfg4"]
frames = [pkg.synth.lidar_frame(c["n"], "kitti", i) for i in range(B)]
cat = np.concatenate(frames)
offs = np.concatenate([[0], np.cumsum([f.shape[0] for f in frames])]).astype(np.int32)
hp = pkg.pinned_empty(cat.shape, np.float32); hp[:] = cat
cap, T = c["max_voxels"], c["max_num_points"]
out = (pkg.pinned_empty((B * cap, T, 4), np.float32), pkg.pinned_empty((B * cap, 4), np.int32),
       pkg.pinned_empty((B * cap,), np.int32), pkg.pinned_empty((B + 1,), np.int32))
kw = {}
if "prefix_upload" in pkg.VoxelGenerator.__init__.__code__.co_varnames:
    kw["prefix_upload"] = (len(sys.argv) > 1 and sys.argv[1] == "prefix")
gen = pkg.VoxelGenerator(c["voxel_size"], c["point_cloud_range"], T, cap, **kw)
for _ in range(3):
    gen.generate_batched(hp, offs, out=out)
t0 = time.perf_counter()
for _ in range(10):
    gen.generate_batched(hp, offs, out=out)
wall = (time.perf_counter() - t0) / 10
gen.profile(True)
for _ in range(10):
    gen.generate_batched(hp, offs, out=out)
pr = gen.profile_read()
print(f"wall={wall*1e3:.3f}ms/call  per launch: " + " ".join(f"{k}={v[0]/v[1]*1e3:.1f}us(x{v[1]//10})" for k, v in pr.items()))
