"""Experiment sweep: per-kernel device times of generate_batched on cfg4-shaped input for the
library's experiment switches (environment variables read when a handle is created)."""
import importlib, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("3dobject_detection_temporal_b200")
frames_n = 64
steps = 30
c = pkg.synth.CONFIGS["cfg4"]
frames = [pkg.synth.lidar_frame(c["n"], "kitti", i) for i in range(frames_n)]
offs = np.concatenate([[0], np.cumsum([f.shape[0] for f in frames])]).astype(np.int32)
sets = [torch.from_numpy(np.concatenate([frames[(i + 7 * k) % frames_n] for i in range(frames_n)])).cuda()
        for k in range(3)]
SW = ["BVX_WARP_TICKETS", "BVX_SMALL_TABLE", "BVX_INSERT_BLOCKS", "BVX_ABLATE"]


def run(env, overflow):
    for k in SW:
        os.environ.pop(k, None)
    os.environ.update(env)
    gen = pkg.VoxelGenerator(c["voxel_size"], c["point_cloud_range"], c["max_num_points"], c["max_voxels"],
                             overflow=overflow)
    out = gen.generate_batched(sets[0], offs, mean_channels=4, padded=True)
    for i in range(5):
        gen.generate_batched(sets[i % 3], offs, mean_channels=4, padded=True, out=out)
    torch.cuda.synchronize()
    best = 1e9
    for rep in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            gen.generate_batched(sets[i % 3], offs, mean_channels=4, padded=True, out=out)
        e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / steps)
    gen.profile(True)
    for i in range(steps):
        gen.generate_batched(sets[i % 3], offs, mean_channels=4, padded=True, out=out)
    torch.cuda.synchronize()
    pr = gen.profile_read()
    chk = int(out[2].sum()) if isinstance(out, (tuple, list)) else -1
    print(f"{overflow:8s} {str(env):50s} step={best*1e3:6.1f}us " +
          " ".join(f"{k}={v[0]/v[1]*1e3:.1f}" for k, v in pr.items() if v[1]) + f" chk={chk}", flush=True)
    gen.close()


for ov in ("break", "continue"):
    run({}, ov)
