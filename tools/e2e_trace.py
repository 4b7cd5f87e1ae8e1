"""End-to-end arm of bench.py only (cfg4, pinned numpy in/out), with BVX_TRACE lines per call on stderr and the
wall time of every call per rank.  usage: [torchrun ...] python tools/e2e_trace.py [steps=10] [host_threads]"""
import importlib, os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("3dobject_detection_temporal_b200")
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 10
threads = int(sys.argv[2]) if len(sys.argv) > 2 else None
rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("WORLD_SIZE", 1), ("LOCAL_RANK", 0)))
torch.cuda.set_device(local)
if world > 1:
    torch.distributed.init_process_group("nccl", device_id=torch.device("cuda", local))
c = pkg.synth.CONFIGS["cfg4"]
B, T, cap = 64, c["max_num_points"], c["max_voxels"]
sets = []
for k in range(3):
    fr = [pkg.synth.lidar_frame(c["n"], "kitti", 1000 * k + g) for g in pkg.shard_frames(B * world, rank, world)]
    hp = pkg.pinned_empty((B * c["n"], 4), np.float32); hp[:] = np.concatenate(fr); sets.append(hp)
offs = (np.arange(B + 1) * c["n"]).astype(np.int32)
out = (pkg.pinned_empty((B * cap, T, 4), np.float32), pkg.pinned_empty((B * cap, 4), np.int32),
       pkg.pinned_empty((B * cap,), np.int32), pkg.pinned_empty((B + 1,), np.int32))
gen = pkg.VoxelGenerator(c["voxel_size"], c["point_cloud_range"], T, cap, device=torch.device("cuda", local),
                         host_threads=threads)
for i in range(3):
    gen.generate_batched(sets[i % 3], offs, out=out)
if world > 1:
    torch.distributed.barrier()
ts = []
for i in range(steps):
    t0 = time.perf_counter(); gen.generate_batched(sets[i % 3], offs, out=out); ts.append((time.perf_counter() - t0) * 1e3)
print(f"rank {rank} pid {os.getpid()} threads {gen._host_threads} calls ms: " + " ".join(f"{t:.2f}" for t in ts), flush=True)
if world > 1:
    torch.distributed.destroy_process_group()
