"""Experiment: the cfg4 batch as S sub-batches on S streams (one handle each), so that the HBM-bound gather of
one sub-batch runs beside the atomic-latency-bound insert of the next.
usage: python tools/split_streams.py [frames=64] [overflow=break] [steps=100] [splits=1,2,4]"""
import importlib, os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
pkg = importlib.import_module("3dobject_detection_temporal_b200")
frames_n = int(sys.argv[1]) if len(sys.argv) > 1 else 64
overflow = sys.argv[2] if len(sys.argv) > 2 else "break"
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 100
splits = [int(s) for s in (sys.argv[4] if len(sys.argv) > 4 else "1,2,4").split(",")]
c = pkg.synth.CONFIGS["cfg4"]
frames = [pkg.synth.lidar_frame(c["n"], "kitti", i) for i in range(frames_n)]
sets = [[torch.from_numpy(frames[(i + 7 * k) % frames_n]).cuda() for i in range(frames_n)] for k in range(3)]
for S in splits:
    per = frames_n // S
    gens = [pkg.VoxelGenerator(c["voxel_size"], c["point_cloud_range"], c["max_num_points"], c["max_voxels"],
                               overflow=overflow) for _ in range(S)]
    streams = [torch.cuda.Stream() for _ in range(S)]
    ins = [[torch.cat(sets[k][s * per:(s + 1) * per]) for s in range(S)] for k in range(3)]
    offs = [np.concatenate([[0], np.cumsum([sets[0][s * per + i].shape[0] for i in range(per)])]).astype(np.int32)
            for s in range(S)]
    outs = []
    for s in range(S):
        with torch.cuda.stream(streams[s]):
            outs.append(gens[s].generate_batched(ins[0][s], offs[s], mean_channels=4, padded=True))
    torch.cuda.synchronize()

    def step(i):
        for s in range(S):
            with torch.cuda.stream(streams[s]):
                gens[s].generate_batched(ins[i % 3][s], offs[s], mean_channels=4, padded=True, out=outs[s])

    for i in range(6):
        step(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    main = torch.cuda.current_stream()
    e0.record(main)
    for s in range(S):
        streams[s].wait_event(e0)
    for i in range(steps):
        step(i)
    for s in range(S):
        main.wait_stream(streams[s])
    e1.record(main)
    torch.cuda.synchronize()
    print(f"splits={S} frames={frames_n} {overflow} step={e0.elapsed_time(e1) / steps * 1e3:.1f}us", flush=True)
    del gens, outs, ins
