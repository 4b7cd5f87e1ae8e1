"""Single-call device times of the other BASELINE configs (cfg1, cfg2, cfg3, cfg5) with CUDA
events: latency of one generate()/generate_sweeps() call on device-resident input, per kernel."""
import importlib, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("3dobject_detection_temporal_b200")
synth = pkg.synth
names = sys.argv[1:] or ["cfg1", "cfg2", "cfg3", "cfg5"]
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for name in names:
    c = synth.CONFIGS[name]
    gen = pkg.VoxelGenerator(c["voxel_size"], c["point_cloud_range"], c["max_num_points"], c["max_voxels"])
    if c["sweeps"] == 1:
        pts = torch.from_numpy(synth.make_frames(name, 1)[0]).cuda()
        call = lambda: gen.generate(pts)
        n, f_out = pts.shape[0], 4
    else:
        sw = [(torch.from_numpy(p).cuda(), R, t, dt) for p, R, t, dt in synth.make_sweeps(name)]
        call = lambda: gen.generate_sweeps(sw)
        n, f_out = sum(s[0].shape[0] for s in sw), 5
    for _ in range(5):
        out = call()
    m = int(out[0].shape[0]); kept = int(out[2].sum())
    T = c["max_num_points"]
    torch.cuda.synchronize()
    ts = []
    for _ in range(20):
        flush.zero_()                      # cold L2, like a frame arriving from the sensor
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); call(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ms = float(np.median(ts))
    gen.profile(True)
    for _ in range(20):
        call()
    torch.cuda.synchronize()
    pr = gen.profile_read()
    # the same call as a CUDA graph replay (padded single-frame form only): no Python / launch overhead
    graph_us = None
    if c["sweeps"] == 1:
        gcall = lambda: gen.generate_padded(pts)
        gcall(); torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            gcall()
        tg = []
        for _ in range(20):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); graph.replay(); e1.record(); torch.cuda.synchronize()
            tg.append(e0.elapsed_time(e1))
        graph_us = float(np.median(tg)) * 1e3
    alg = 4 * n * 4 + m * (T * f_out * 4 + 12 + 4)
    print(f"{name}: points={n} voxels={m} kept={kept} T={T} call={ms*1e3:.1f}us -> {n/ms/1e3:.0f} Mpoints/s, "
          f"algorithmic {alg/1e6:.1f} MB -> {alg/ms/1e6:.0f} GB/s" +
          (f", graph replay {graph_us:.1f}us" if graph_us else "") + " | " +
          " ".join(f"{k}={v[0]/v[1]*1e3:.1f}" for k, v in pr.items() if v[1]))
    gen.close()
# fused pillar path (cfg3): points -> canvas in one go vs generate + mean VFE + scatter
c = synth.CONFIGS["cfg3"]
gen = pkg.VoxelGenerator(c["voxel_size"], c["point_cloud_range"], c["max_num_points"], c["max_voxels"])
pts = torch.from_numpy(synth.make_frames("cfg3", 1)[0]).cuda()
offs = np.asarray([0, pts.shape[0]], np.int32)
vfe, gx, gy, gz = pkg.VoxelFeatureExtractorV3(), *[int(x) for x in gen.grid_size]
sc = pkg.PointPillarsScatter([1, gz, gy, gx, 4])
out = gen.generate_bev(pts, offs)
def fused(): gen.generate_bev(pts, offs, out=out)
def separate():
    v, cc, n, off = gen.generate_batched(pts, offs)
    sc(vfe(v, n), cc, 1)
for nm, fn in (("fused generate_bev", fused), ("generate_batched + VFE + scatter", separate)):
    for _ in range(5): fn()
    torch.cuda.synchronize(); ts = []
    for _ in range(20):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    print(f"cfg3 points -> canvas, {nm}: {float(np.median(ts))*1e3:.1f}us")

