"""Where the pipelined host form spends its time (BVX_TRACE=1 prints the legs of every call)."""
import importlib, os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("3dobject_detection_temporal_b200")
c = pkg.synth.CONFIGS["cfg4"]
B, T, F, cap = 64, 5, 4, c["max_voxels"]
frames = [pkg.synth.lidar_frame(c["n"], "kitti", i) for i in range(B)]
offs = np.concatenate([[0], np.cumsum([f.shape[0] for f in frames])]).astype(np.int32)
host = np.concatenate(frames, axis=0)
hp = pkg.pinned_empty(host.shape, np.float32); hp[:] = host
hout = (pkg.pinned_empty((B * cap, T, F), np.float32), pkg.pinned_empty((B * cap, 4), np.int32),
        pkg.pinned_empty((B * cap,), np.int32), pkg.pinned_empty((B + 1,), np.int32))
for asm, mod in ((True, 0), (False, 0)):
    g = pkg.VoxelGenerator(c["voxel_size"], c["point_cloud_range"], T, cap, host_assemble=asm)
    for _ in range(3):
        g.generate_batched(hp, offs, out=hout)
    os.environ["BVX_TRACE"] = "1"
    t0 = time.perf_counter()
    for _ in range(3):
        g.generate_batched(hp, offs, out=hout)
    print(f"assemble={asm}: {(time.perf_counter() - t0) / 3 * 1e3:.3f} ms/call", flush=True)
    os.environ.pop("BVX_TRACE")
    g.close()
