"""Device timeline of one steady-state step (experiment build with -DBVX_TIMELINE): first block start / last block
end of k_insert's clear rows and chunk rows, k_mid, k_gather, relative to the step's first event.
usage: BVX_LIB=.../libbvx_tl.so python tools/timeline.py [frames=64] [overflow=break]"""
import ctypes, importlib, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("3dobject_detection_temporal_b200")
frames_n = int(sys.argv[1]) if len(sys.argv) > 1 else 64
overflow = sys.argv[2] if len(sys.argv) > 2 else "break"
c = pkg.synth.CONFIGS["cfg4"]
frames = [pkg.synth.lidar_frame(c["n"], "kitti", i) for i in range(frames_n)]
offs = np.concatenate([[0], np.cumsum([f.shape[0] for f in frames])]).astype(np.int32)
sets = [torch.from_numpy(np.concatenate([frames[(i + 7 * k) % frames_n] for i in range(frames_n)])).cuda() for k in range(3)]
gen = pkg.VoxelGenerator(c["voxel_size"], c["point_cloud_range"], c["max_num_points"], c["max_voxels"], overflow=overflow)
out = gen.generate_batched(sets[0], offs, mean_channels=4, padded=True)
lib = ctypes.CDLL(pkg._lib.LIB_PATH)
buf = (ctypes.c_uint64 * 32)()
names = ["insert.clear", "insert.chunks", "mid", "gather", "mid.scanned", "mid.numbered", "mid.published", "mid.started", "mid.loaded"]
for i in range(6):
    gen.generate_batched(sets[i % 3], offs, mean_channels=4, padded=True, out=out)
lib.bvx_debug_timeline(buf, 1)
NS = 3
for i in range(NS):
    gen.generate_batched(sets[i % 3], offs, mean_channels=4, padded=True, out=out)
    if i < NS - 1:
        pass
# only the last step's stamps are clean if we reset right before it: do one isolated-in-stream step between two resets
lib.bvx_debug_timeline(buf, 1)
for i in range(2):            # two back-to-back steps: min = first step's starts, max = second step's ends
    gen.generate_batched(sets[i % 3], offs, mean_channels=4, padded=True, out=out)
lib.bvx_debug_timeline(buf, 1)
t0 = min(buf[2 * k] for k in range(9) if buf[2 * k] != 2**64 - 1)
print(f"frames={frames_n} {overflow}: two back-to-back steps, us from first event (start of first = step 1, end of last = step 2)")
for k, n in enumerate(names):
    print(f"  {n:14s} first start {(buf[2*k]-t0)/1e3:8.1f}   last end {(buf[2*k+1]-t0)/1e3:8.1f}")
gen.generate_batched(sets[0], offs, mean_channels=4, padded=True, out=out); torch.cuda.synchronize()
lib.bvx_debug_timeline(buf, 1)
gen.generate_batched(sets[1], offs, mean_channels=4, padded=True, out=out)
lib.bvx_debug_timeline(buf, 1)
t0 = min(buf[2 * k] for k in range(9) if buf[2 * k] != 2**64 - 1)
print("  one isolated step:")
for k, n in enumerate(names):
    print(f"  {n:14s} first start {(buf[2*k]-t0)/1e3:8.1f}   last end {(buf[2*k+1]-t0)/1e3:8.1f}")
