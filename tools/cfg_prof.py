"""Per-kernel device times of one BASELINE config through bench.py's Arm (experiments).
usage: python tools/cfg_prof.py cfg2 [frames=1] [overflow=break] [steps=30]"""
import argparse, importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
cfg = sys.argv[1]
frames = int(sys.argv[2]) if len(sys.argv) > 2 else 1
overflow = sys.argv[3] if len(sys.argv) > 3 else "break"
steps = int(sys.argv[4]) if len(sys.argv) > 4 else 30
args = argparse.Namespace(config=cfg, frames=frames, overflow=overflow, scaling="weak", steps=steps)
pkg = importlib.import_module(bench.PKG)
dev = torch.device("cuda", 0)
arm = bench.Arm(args, pkg, torch, dev, 0, 1)
arm.step(0)
m = arm.voxels()
ms, launches = arm.time_steps(steps, 3)
pk = arm.per_kernel(steps)
ab = sum(arm.bytes(m).values())
print(f"{cfg} frames={frames} {overflow} step={ms*1e3:.1f}us voxels={m} points={arm.npts} contract={ab/1e6:.1f}MB "
      f"frac={ab/(ms*1e-3)/1e9/6550:.3f} " + " ".join(f"{k}={v*1e3:.1f}" for k, v in pk.items()), flush=True)
