#!/usr/bin/env python
"""bench.py -- headline benchmark of the voxelization hot path.

Metric (BASELINE.json): voxelized Mpoints/s (input points, counted before range
filtering), with the fraction of the HBM roofline of the dominant kernel and the
CPU oracle timed on this box's host cores.

Workload at every N: BASELINE cfg4 -- 64 synthetic KITTI-shaped 120k-point frames
PER GPU (frames shard by index, weak scaling, no collective on the data path),
SECOND car voxelizer 0.05 x 0.05 x 0.1 m, T=5, max_voxels=20000, overflow=break,
mean VFE fused into the gather.  A "step" = one pass of generate_batched over the
rank's 64 frames.

    python bench.py --gpus N --steps K --warmup W          (our arm)
    python bench.py --impl reference ...                   (CPU oracle, all host threads)
"""
import argparse
import importlib
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
PKG = "3dobject_detection_temporal_b200"

METRIC = "voxelized_points_per_second"
UNIT = "Mpoints/s"
WORKLOAD = ("cfg4: 64 KITTI-shaped 120k-point frames per GPU, SECOND car voxelizer "
            "0.05x0.05x0.1 m, T=5, max_voxels=20000, overflow=break, hashed grid, fused mean VFE")


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--frames", type=int, default=64, help="frames per GPU")
    ap.add_argument("--overflow", default="break", choices=["break", "continue"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-numa-bind", action="store_true", help="do not pin the process to the GPU's NUMA node")
    return ap.parse_args()


def measured_peaks():
    """(hbm_gbs, source) -- MEASURED_PEAKS.json if the driver wrote it, else the
    fallback B200_PROFILING.md states (6.65 TB/s)."""
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index=0):
        self.gpu = gpu_index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                 "-i", str(self.gpu)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            c = [x.strip() for x in r.split(",")]
            if len(c) < 9:
                continue
            try:
                sm.append(float(c[1])); mx.append(float(c[2]))
            except ValueError:
                continue
            for k, nm in enumerate(names):
                if c[5 + k].lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": statistics.median(sm) if sm else None,
                "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


# --------------------------------------------------------------------- workload
def build_frames(synth, rank, world, frames_per_gpu):
    """The rank's shard of the 64*world-frame job: global frame g -> rank g % world
    (DistributedSampler interleave), seed = global frame index."""
    pkg = importlib.import_module(PKG)
    idx = pkg.shard_frames(frames_per_gpu * world, rank, world)
    c = synth.CONFIGS["cfg4"]
    return [synth.lidar_frame(c["n"], c["sensor"], c["seed0"] + g) for g in idx], c


def algorithmic_bytes(n_points, f, m_voxels, t, mean_c):
    """generate = 4*(N*F_in + M*T*F_out + 4*M (batched coords) + M);
    fused mean VFE adds its compulsory output 4*M*C."""
    return dict(points_in=4 * n_points * f,
                voxels_out=4 * m_voxels * t * f,
                coors_out=16 * m_voxels,
                num_out=4 * m_voxels,
                mean_out=4 * m_voxels * mean_c)


# ------------------------------------------------------------------ CPU baseline
def cpu_oracle_single_core(frames, c, overflow, budget_s=12.0):
    """Variant A of BASELINE.md section 4: the C oracle (spconv C++ twin), 1 core, kernel
    only -- dense grid cached and reset, outputs pre-allocated."""
    import oracle
    g = oracle.OracleVoxelGenerator(c["voxel_size"], c["point_cloud_range"], c["max_num_points"],
                                    c["max_voxels"], overflow)
    t_, f = c["max_num_points"], frames[0].shape[1]
    voxels = np.zeros((c["max_voxels"], t_, f), np.float32)
    coors = np.zeros((c["max_voxels"], 3), np.int32)
    num = np.zeros((c["max_voxels"],), np.int32)
    g.kernel(frames[0], c["max_voxels"], voxels, coors, num)        # warm: page in the grid
    done, t_kernel = 0, 0.0
    t_start = time.perf_counter()
    while time.perf_counter() - t_start < budget_s:
        for fr in frames:
            voxels.fill(0); num.fill(0)                              # the reference zero-allocates
            t0 = time.perf_counter()
            g.kernel(fr, c["max_voxels"], voxels, coors, num)
            t_kernel += time.perf_counter() - t0
            done += fr.shape[0]
        if done >= 40 * sum(fr.shape[0] for fr in frames):
            break
    return done / t_kernel / 1e6, done


def cpu_oracle_all_threads(frames, c, overflow, steps, warmup):
    """The reference's own parallelisation model: P workers, one frame each
    (num_workers in second/pytorch/train.py:259-266).  P threads over the C oracle
    (ctypes releases the GIL), each with its own cached dense grid."""
    import concurrent.futures as cf
    import oracle
    import psutil
    cores = os.cpu_count() or 1
    grid_bytes = int(np.prod(oracle.grid_size(c["voxel_size"], c["point_cloud_range"]))) * 4
    mem_ok = max(1, int(psutil.virtual_memory().available * 0.5 // (grid_bytes + (8 << 20))))
    p = max(1, min(cores, len(frames), mem_ok))
    gens = [oracle.OracleVoxelGenerator(c["voxel_size"], c["point_cloud_range"], c["max_num_points"],
                                        c["max_voxels"], overflow) for _ in range(p)]
    t_, f = c["max_num_points"], frames[0].shape[1]
    bufs = [(np.zeros((c["max_voxels"], t_, f), np.float32), np.zeros((c["max_voxels"], 3), np.int32),
             np.zeros((c["max_voxels"],), np.int32)) for _ in range(p)]

    def work(w):
        g, (v, co, nu) = gens[w], bufs[w]
        for k in range(w, len(frames), p):
            v.fill(0); nu.fill(0)
            g.kernel(frames[k], c["max_voxels"], v, co, nu)

    with cf.ThreadPoolExecutor(p) as ex:
        for _ in range(max(warmup, 1)):
            list(ex.map(work, range(p)))
        t0 = time.perf_counter()
        for _ in range(steps):
            list(ex.map(work, range(p)))
        dt = time.perf_counter() - t0
    return dt / steps, p


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    pkg = importlib.import_module(PKG)
    frames, c = build_frames(pkg.synth, 0, 1, args.frames)
    npts = sum(f.shape[0] for f in frames)
    steps = max(1, min(args.steps, 20))
    sec, p = cpu_oracle_all_threads(frames, c, args.overflow, steps, min(args.warmup, 2))
    val = npts / sec / 1e6
    sample = f"{len(frames)} frames x {c['n']} points per step, {steps} steps, C oracle, one frame per thread"
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
            "steps": steps, "warmup": min(args.warmup, 2), "ms_per_step": sec * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": WORKLOAD, "overflow": args.overflow,
                       "note": "CPU restatement of the reference voxelizer (spconv is not in the reference "
                               "tree); the reference arm does not scale with GPUs"},
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": p, "kind": "port", "sample": sample},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))
    return 0


# ----------------------------------------------------------------------- ours
def run_ours(args):
    import torch
    import torch.distributed as dist
    pkg = importlib.import_module(PKG)
    synth = pkg.synth
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device: the product path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    numa = None if args.no_numa_bind else pkg.shard.bind_host_to_device(local)   # before any pinned allocation
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    frames, c = build_frames(synth, rank, world, args.frames)
    B, T, F, cap, C = len(frames), c["max_num_points"], 4, c["max_voxels"], 4
    offs = np.concatenate([[0], np.cumsum([f.shape[0] for f in frames])]).astype(np.int32)
    npts = int(offs[-1])
    host = np.concatenate(frames, axis=0)
    # L2 hygiene: the timed steps rotate over NSETS distinct input batches (each 123 MB,
    # together > the 126 MB L2); every step also writes ~128 MB of outputs.
    NSETS = 3
    sets = []
    for k in range(NSETS):
        order = np.roll(np.arange(B), -7 * k)
        sets.append(torch.from_numpy(np.concatenate([frames[i] for i in order], axis=0)).to(dev))
    gen = pkg.VoxelGenerator(c["voxel_size"], c["point_cloud_range"], T, cap, overflow=args.overflow, device=dev)
    out = gen.generate_batched(sets[0], offs, mean_channels=C, padded=True)
    torch.cuda.synchronize()
    m_voxels = int(out[3][-1].item())
    kept = int(out[2][:m_voxels].sum().item())

    def step(i):
        gen.generate_batched(sets[i % NSETS], offs, mean_channels=C, padded=True, out=out)

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    for i in range(max(args.warmup, 3)):
        step(i)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    l0 = pkg._lib.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.steps):
        step(i)
    e1.record()
    torch.cuda.synchronize()
    launches = pkg._lib.launch_count() - l0
    ms_total = e0.elapsed_time(e1)
    if world > 1:
        dist.barrier()
        t = torch.tensor([ms_total], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total = float(t.item())
    ms_step = ms_total / args.steps
    total_pts, total_vox = pkg.shard.gather_counts(npts, m_voxels) if world > 1 else (npts, m_voxels)
    value = total_pts / (ms_step * 1e-3) / 1e6

    # ---- per-kernel device times (CUDA events on the launching stream), same steps
    gen.profile(True)
    for i in range(args.steps):
        step(i)
    torch.cuda.synchronize()
    prof = gen.profile_read()
    gen.profile(False)
    per_kernel = {k: v[0] / max(v[1], 1) for k, v in prof.items()}          # ms per launch
    ksum = sum(per_kernel.values())
    dom = max(per_kernel, key=per_kernel.get)
    ab = algorithmic_bytes(npts, F, m_voxels, T, C)
    # algorithmic bytes of each kernel = its share of the contract's compulsory I/O
    k_bytes = {"k_insert": ab["points_in"],
               "k_gather": ab["voxels_out"] + ab["coors_out"] + ab["num_out"] + ab["mean_out"]}
    peak, peak_src = measured_peaks()
    dom_bytes = k_bytes.get(dom, 0)
    achieved = dom_bytes / (per_kernel[dom] * 1e-3) / 1e9
    pipe_bytes = sum(ab.values())
    pipe_gbs = pipe_bytes / (ms_step * 1e-3) / 1e9
    # dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed ncu --set full
    # capture of this workload (profiles/r01_f_ncu_full_summary.txt); cold L2, cfg4 with 64 frames only
    ncu_traffic = {"k_insert": 66.47e6 + 66.16e6, "k_verify": 1.05e6 + 0.00e6,
                   "k_slot": 30.41e6 + 4.14e6, "k_gather": 74.95e6 + 96.87e6}
    traffic = ncu_traffic.get(dom) if (B == 64 and args.overflow == "break") else None
    roofline = {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": traffic,
                "traffic_source": "profiles/r01_f_ncu_full_summary.txt (ncu --set full, cold L2)",
                "peak_source": peak_src,
                "algorithmic_bytes_per_launch": dom_bytes,
                "kernel_ms": per_kernel[dom], "kernel_share_of_step": per_kernel[dom] / ksum,
                "per_kernel_ms": per_kernel,
                "pipeline": {"algorithmic_bytes_per_step": pipe_bytes, "achieved": pipe_gbs,
                             "frac": pipe_gbs / peak}}

    # ---- end to end through the host-buffer call: pinned numpy in, pinned numpy out
    e2e = e2e_full = e2e_rows = None
    if not args.no_e2e:
        hp = pkg.pinned_empty(host.shape, np.float32)
        hp[:] = host
        hout = (pkg.pinned_empty((B * cap, T, F), np.float32), pkg.pinned_empty((B * cap, 4), np.int32),
                pkg.pinned_empty((B * cap,), np.int32), pkg.pinned_empty((B + 1,), np.int32))
        ksteps = max(3, min(args.steps, 20))

        def host_arm(prefix_upload, host_assemble=True):
            gen_h = pkg.VoxelGenerator(c["voxel_size"], c["point_cloud_range"], T, cap, overflow=args.overflow,
                                       device=dev, prefix_upload=prefix_upload, host_assemble=host_assemble)
            gen_threads = gen_h._host_threads or min(32, len(os.sched_getaffinity(0)))
            for _ in range(3):
                r = gen_h.generate_batched(hp, offs, out=hout)
            if world > 1:
                dist.barrier()
            t0 = time.perf_counter()
            h2d = d2h = redone = 0
            for _ in range(ksteps):
                r = gen_h.generate_batched(hp, offs, out=hout)      # returns after the D2H completed
                a, b_, rd = gen_h.last_transfer()                   # bytes the library actually copied
                h2d += a; d2h += b_; redone += rd
            dt = (time.perf_counter() - t0) / ksteps
            if world > 1:
                t = torch.tensor([dt], dtype=torch.float64, device=dev)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                dt = float(t.item())
            gen_h.close()
            return {"value": total_pts / dt / 1e6, "unit": UNIT, "ms_per_step": dt * 1e3, "steps": ksteps,
                    "h2d_bytes_per_step": h2d // ksteps, "d2h_bytes_per_step": d2h // ksteps,
                    "sub_batches_redone": redone,
                    "api": "VoxelGenerator.generate_batched(numpy) -> bvx_generate_batched_host; pinned host "
                           "buffers" + ("; break mode uploads only the prefix of each frame the cut-off can reach "
                                        "(sized from the previous call, verified per frame, redone if short)"
                                        if prefix_upload and args.overflow == "break" else "; whole frames uploaded")
                           + ("; the kept rows come home compact (16 bytes per kept point), the zero padding is "
                              "written by the handle's host threads" if host_assemble else
                              "; padded rows gathered on the device and copied home"),
                    "host_threads": (gen_threads if host_assemble else 0),
                    "host_numa_binding": numa}

        e2e = host_arm(True)
        e2e_rows = host_arm(True, host_assemble=False)
        if args.overflow == "break":
            e2e_full = host_arm(False, host_assemble=False)
    clocks = sampler.stop() if rank == 0 else None

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        v, done = cpu_oracle_single_core(frames, c, args.overflow)
        cpu = {"value": v, "unit": UNIT, "cores": 1, "kind": "port",
               "sample": f"{done} points ({done // npts} passes over the same {B} frames), C oracle kernel only "
                         "(dense grid cached and reset, outputs pre-allocated)"}
    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": WORKLOAD, "frames_per_gpu": B, "points_per_gpu": npts,
                           "voxels_per_gpu": m_voxels, "kept_points_per_gpu": kept, "overflow": args.overflow,
                           "l2": f"inputs rotate over {NSETS} distinct {host.nbytes >> 20} MiB batches "
                                 "(> 126 MB L2); outputs rewritten every step; no explicit flush"},
                "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "e2e_device_rows": e2e_rows, "e2e_full_upload": e2e_full,
                "gpu_launches": int(launches),
                "clocks": clocks}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    args = parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
