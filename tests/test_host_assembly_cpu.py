"""Host half of the pipelined host form (no GPU): the padded [M, T, f] rows assembled by the
library's host threads from point indices / compact kept rows must equal a plain numpy
restatement -- ragged frames, empty frames, full and single-point voxels, both the 16-byte
streaming path and the generic path, one and many threads."""
import ctypes
import importlib

import numpy as np
import pytest

pkg = importlib.import_module("3dobject_detection_temporal_b200")


def _ptr(a):
    return ctypes.c_void_p(a.ctypes.data)


def _aligned(shape, dtype, offset_bytes=0):
    n = int(np.prod(shape)) * np.dtype(dtype).itemsize
    raw = np.zeros(n + 64, np.uint8)
    o = (-raw.ctypes.data) % 64 + offset_bytes
    return raw[o:o + n].view(dtype).reshape(shape)


def _structure(rng, frame_points, voxels_per_frame, T):
    """Random voxel structure: per frame M_i voxels, each with 1..T distinct point indices."""
    offs = np.concatenate([[0], np.cumsum(frame_points)]).astype(np.int32)
    fvo = np.concatenate([[0], np.cumsum(voxels_per_frame)]).astype(np.int32)
    m = int(fvo[-1])
    num = np.zeros(m, np.int32)
    idx = np.full((m, T), -1, np.int32)
    for i, (n, mv) in enumerate(zip(frame_points, voxels_per_frame)):
        for v in range(fvo[i], fvo[i] + mv):
            k = int(rng.integers(1, min(T, n) + 1))
            if v % 7 == 0:
                k = min(T, n)
            num[v] = k
            idx[v, :k] = rng.choice(n, size=k, replace=False)
    return offs, fvo, num, idx


def _want(points, offs, fvo, num, idx, T, f):
    m = int(fvo[-1])
    out = np.zeros((m, T, f), np.float32)
    for i in range(len(fvo) - 1):
        for v in range(fvo[i], fvo[i + 1]):
            out[v, :num[v]] = points[offs[i] + idx[v, :num[v]]]
    return out


@pytest.mark.parametrize("T,f,threads,misalign", [(5, 4, 1, 0), (5, 4, 7, 0), (3, 4, 4, 4), (10, 5, 3, 0), (1, 7, 16, 0)])
def test_assemble_rows(T, f, threads, misalign):
    lib = pkg._lib.load()
    rng = np.random.default_rng(T * 31 + f)
    frame_points = [900, 0, 40, 1500, 12]
    vpf = [300, 0, 17, 450, 3]
    offs, fvo, num, idx = _structure(rng, frame_points, vpf, T)
    points = _aligned((int(offs[-1]), f), np.float32, misalign)
    points[:] = rng.normal(0, 30, size=points.shape).astype(np.float32)
    m = int(fvo[-1])
    out = _aligned((m, T, f), np.float32)
    out[:] = 7.0
    rc = lib.bvx_host_assemble_rows(_ptr(points), _ptr(offs), _ptr(fvo), len(frame_points), T, f, _ptr(num),
                                    _ptr(idx), _ptr(out), threads)
    assert rc == 0
    assert np.array_equal(out.view(np.uint32), _want(points, offs, fvo, num, idx, T, f).view(np.uint32))


@pytest.mark.parametrize("T,threads,cap", [(5, 1, 700), (5, 6, 700), (8, 3, 300), (2, 32, 256)])
def test_assemble_compact(T, threads, cap):
    """Compact rows as k_gather_small writes them: tiles of 256 voxels in arbitrary (arrival) order."""
    lib = pkg._lib.load()
    rng = np.random.default_rng(T + cap)
    frame_points = [4000, 0, 9, 2500]
    vpf = [cap, 0, 5, min(cap, 333)]
    offs, fvo, num, idx = _structure(rng, frame_points, vpf, T)
    points = rng.normal(0, 30, size=(int(offs[-1]), 4)).astype(np.float32)
    tiles = (cap + 255) // 256
    b = len(frame_points)
    # hand the tiles their ranges in a shuffled order, like the atomic counter does
    tile_rows = np.zeros((b, tiles), np.int64)
    for i in range(b):
        for t in range(tiles):
            lo, hi = fvo[i] + t * 256, min(fvo[i + 1], fvo[i] + t * 256 + 256)
            tile_rows[i, t] = int(num[lo:hi].sum()) if hi > lo else 0
    order = rng.permutation(b * tiles)
    base = np.zeros(b * tiles, np.int32)
    run = 0
    for j in order:
        base[j] = run
        run += int(tile_rows.reshape(-1)[j])
    rows = _aligned((max(run, 1), 4), np.float32)
    for i in range(b):
        for t in range(tiles):
            r = int(base[i * tiles + t])
            for v in range(fvo[i] + t * 256, min(fvo[i + 1], fvo[i] + t * 256 + 256)):
                rows[r:r + num[v]] = points[offs[i] + idx[v, :num[v]]]
                r += int(num[v])
    m = int(fvo[-1])
    out = _aligned((m, T, 4), np.float32)
    out[:] = 7.0
    rc = lib.bvx_host_assemble_compact(_ptr(rows), _ptr(base), _ptr(fvo), b, cap, T, _ptr(num), _ptr(out), threads)
    assert rc == 0
    assert np.array_equal(out.view(np.uint32), _want(points, offs, fvo, num, idx, T, 4).view(np.uint32))


def test_assemble_rejects_bad_arguments():
    lib = pkg._lib.load()
    z = np.zeros(4, np.int32)
    assert lib.bvx_host_assemble_rows(None, _ptr(z), _ptr(z), 0, 5, 4, _ptr(z), _ptr(z), _ptr(z), 1) != 0
    rows = _aligned((4, 4), np.float32, 4)          # misaligned
    out = _aligned((1, 5, 4), np.float32)
    assert lib.bvx_host_assemble_compact(_ptr(rows), _ptr(z), _ptr(z), 1, 256, 5, _ptr(z), _ptr(out), 1) != 0
