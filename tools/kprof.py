"""Per-kernel device times of generate_batched on cfg4-shaped input (experiments).
usage: python tools/kprof.py [frames=64] [overflow=break] [steps=30]"""
import importlib, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("3dobject_detection_temporal_b200")
frames_n = int(sys.argv[1]) if len(sys.argv) > 1 else 64
overflow = sys.argv[2] if len(sys.argv) > 2 else "break"
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 30
c = pkg.synth.CONFIGS["cfg4"]
frames = [pkg.synth.lidar_frame(c["n"], "kitti", i) for i in range(frames_n)]
offs = np.concatenate([[0], np.cumsum([f.shape[0] for f in frames])]).astype(np.int32)
sets = [torch.from_numpy(np.concatenate([frames[(i + 7 * k) % frames_n] for i in range(frames_n)])).cuda()
        for k in range(3)]
gen = pkg.VoxelGenerator(c["voxel_size"], c["point_cloud_range"], c["max_num_points"], c["max_voxels"],
                         overflow=overflow, grid=os.environ.get("BVX_GRID", "hashed"))
out = gen.generate_batched(sets[0], offs, mean_channels=4, padded=True)
for i in range(5):
    gen.generate_batched(sets[i % 3], offs, mean_channels=4, padded=True, out=out)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for i in range(steps):
    gen.generate_batched(sets[i % 3], offs, mean_channels=4, padded=True, out=out)
e1.record(); torch.cuda.synchronize()
tot = e0.elapsed_time(e1) / steps
gen.profile(True)
for i in range(steps):
    gen.generate_batched(sets[i % 3], offs, mean_channels=4, padded=True, out=out)
torch.cuda.synchronize()
pr = gen.profile_read()
print(f"{os.environ.get('BVX_LIB', 'libbvx.so').split('/')[-1]} frames={frames_n} {overflow} step={tot*1e3:.1f}us " +
      " ".join(f"{k}={v[0]/v[1]*1e3:.1f}" for k, v in pr.items()), flush=True)
