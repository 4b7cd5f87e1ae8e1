"""End-to-end arm only (cfg4, pinned numpy in/out, three rotating input batches): ms per call.
usage: python tools/e2e_prof.py [steps=20]   (BVX_NO_ENQUEUE_THREAD=1 / BVX_TRACE=1 to compare / trace)"""
import importlib, os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("3dobject_detection_temporal_b200")
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 20
c = pkg.synth.CONFIGS["cfg4"]
B, T, cap = 64, c["max_num_points"], c["max_voxels"]
sets = []
for k in range(3):
    hp = pkg.pinned_empty((B * c["n"], 4), np.float32)
    hp[:] = np.concatenate([pkg.synth.lidar_frame(c["n"], "kitti", 1000 * k + g) for g in range(B)])
    sets.append(hp)
offs = (np.arange(B + 1) * c["n"]).astype(np.int32)
out = (pkg.pinned_empty((B * cap, T, 4), np.float32), pkg.pinned_empty((B * cap, 4), np.int32),
       pkg.pinned_empty((B * cap,), np.int32), pkg.pinned_empty((B + 1,), np.int32))
for pu, ha in ((True, True), (False, True), (True, False)):
    gen = pkg.VoxelGenerator(c["voxel_size"], c["point_cloud_range"], T, cap, prefix_upload=pu, host_assemble=ha)
    for i in range(3):
        gen.generate_batched(sets[i % 3], offs, out=out)
    t0 = time.perf_counter()
    for i in range(steps):
        gen.generate_batched(sets[i % 3], offs, out=out)
    dt = (time.perf_counter() - t0) / steps
    print(f"prefix_upload={pu} host_assemble={ha}: {dt*1e3:.2f} ms/call = {B*c['n']/dt/1e6:.0f} Mpoints/s, transfer {gen.last_transfer()}", flush=True)
    gen.close()
