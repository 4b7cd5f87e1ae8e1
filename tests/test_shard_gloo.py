"""N > 1 path on CPU: two gloo ranks shard a batch of frames exactly like bench.py does
(frame g -> rank g % world, no collective on the data path), each voxelizes its own frames
with the oracle (the GPU is not needed for the host-side logic), and the bookkeeping
collective (sum of points / voxels, max of the elapsed time) gives the whole-job numbers."""
import importlib
import os
import socket
import sys

import numpy as np
import pytest

torch = pytest.importorskip("torch")

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = "3dobject_detection_temporal_b200"


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, nframes, out_dir):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    pkg = importlib.import_module(PKG)
    import oracle
    idx = pkg.shard_frames(nframes, rank, world)
    frames = [pkg.synth.lidar_frame(3000 + 100 * g, "kitti", seed=g) for g in idx]
    gen = oracle.OracleVoxelGenerator([0.05, 0.05, 0.1], [0, -40, -3, 70.4, 40, 1], 5, 800)
    v, c, n, off = gen.generate_batched(frames) if frames else (np.zeros((0, 5, 4), np.float32),) * 4
    npts = sum(f.shape[0] for f in frames)
    tot_pts, tot_vox = pkg.shard.gather_counts(npts, int(v.shape[0]))
    # max-over-ranks of the (fake) elapsed time, as bench.py does with the CUDA-event time
    t = torch.tensor([10.0 + rank], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    np.savez(os.path.join(out_dir, f"rank{rank}.npz"), idx=np.asarray(idx), coors=c, num=n, off=off,
             tot=np.asarray([tot_pts, tot_vox]), tmax=t.numpy())
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.timeout(240)
def test_two_rank_frame_sharding(tmp_path, oracle, synth, pkg):
    import torch.multiprocessing as mp
    world, nframes = 2, 7
    port = _free_port()
    mp.spawn(_worker, args=(world, port, nframes, str(tmp_path)), nprocs=world, join=True)
    res = [np.load(tmp_path / f"rank{r}.npz") for r in range(world)]
    # every frame is owned by exactly one rank, interleaved like DistributedSampler
    owned = sorted(int(i) for r in res for i in r["idx"])
    assert owned == list(range(nframes))
    assert res[0]["idx"].tolist() == [0, 2, 4, 6] and res[1]["idx"].tolist() == [1, 3, 5]
    # whole-job counts from the bookkeeping collective == serial run over all frames
    gen = oracle.OracleVoxelGenerator([0.05, 0.05, 0.1], [0, -40, -3, 70.4, 40, 1], 5, 800)
    frames = [synth.lidar_frame(3000 + 100 * g, "kitti", seed=g) for g in range(nframes)]
    per_frame = [gen.generate(f) for f in frames]
    want_pts = sum(f.shape[0] for f in frames)
    want_vox = sum(v.shape[0] for v, _, _ in per_frame)
    for r in res:
        assert r["tot"].tolist() == [want_pts, want_vox]
        assert float(r["tmax"][0]) == 11.0
    # each rank's results are its own frames' results, untouched by the other rank
    for r in res:
        off = r["off"]
        for j, g in enumerate(r["idx"]):
            sl = slice(off[j], off[j + 1])
            assert np.array_equal(r["coors"][sl, 1:], per_frame[g][1])
            assert np.all(r["coors"][sl, 0] == j)
            assert np.array_equal(r["num"][sl], per_frame[g][2])
