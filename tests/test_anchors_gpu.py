"""GPU parity of the anchors-area path through the C ABI, against
the oracle's restatement of second/core/box_np_ops.py:776-810 and the golden vectors made by
running the reference's own functions (tests/golden/make_golden_anchors.py).
Bar: bit-exact (every value is a voxel count)."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")

KITTI = dict(voxel_size=[0.05, 0.05, 0.1], point_cloud_range=[0, -40, -3, 70.4, 40, 1])


def second_anchors_bv(rg, fm_xy=(176, 200), sizes=((1.6, 3.9),)):
    """Bird's-eye boxes of a SECOND-style anchor grid: one centre per feature-map cell, each size
    at rotations 0 and pi/2 (axis-aligned extents), rows (x0, y0, x1, y1)."""
    xs = np.linspace(rg[0], rg[3], fm_xy[0], dtype=np.float32)
    ys = np.linspace(rg[1], rg[4], fm_xy[1], dtype=np.float32)
    cx, cy = np.meshgrid(xs, ys)
    out = []
    for w, l in sizes:
        for dx, dy in ((l, w), (w, l)):
            out.append(np.stack([cx - dx / 2, cy - dy / 2, cx + dx / 2, cy + dy / 2], -1).reshape(-1, 4))
    return np.ascontiguousarray(np.stack(out, 1).reshape(-1, 4), dtype=np.float32)


def test_golden_vectors(pkg, golden_dir):
    z = np.load(os.path.join(golden_dir, "anchors_reference.npz"))
    for tag in ("pillar", "second", "odd_nx"):
        grid = z[f"{tag}_grid"]
        coors = torch.from_numpy(z[f"{tag}_coors"]).cuda()
        got = pkg.anchors_area(coors, z[f"{tag}_anchors_bv"], z[f"{tag}_voxel_size"], z[f"{tag}_range"], grid)
        assert got.dtype == torch.float32 and np.array_equal(got.cpu().numpy(), z[f"{tag}_area"]), tag
        # the reference's decomposition under its own names, cumsums by the caller
        dense = pkg.sparse_sum_for_anchors_mask(coors, tuple(grid[::-1][1:]))
        want = np.zeros(tuple(grid[::-1][1:]), np.float32)
        nz = z[f"{tag}_hist_nonzero"]
        want[nz[:, 0], nz[:, 1]] = z[f"{tag}_hist_values"]
        assert np.array_equal(dense.cpu().numpy(), want), tag
        dense = dense.cumsum(0).cumsum(1)
        got2 = pkg.fused_get_anchors_area(dense, z[f"{tag}_anchors_bv"], z[f"{tag}_voxel_size"], z[f"{tag}_range"],
                                          grid)
        assert np.array_equal(got2.cpu().numpy(), z[f"{tag}_area"]), tag
        mask = pkg.anchors_mask(coors, z[f"{tag}_anchors_bv"], z[f"{tag}_voxel_size"], z[f"{tag}_range"], grid)
        assert mask.dtype == torch.bool and np.array_equal(mask.cpu().numpy(), z[f"{tag}_area"] > 1)


def test_kitti_frame_from_generate(pkg, oracle, synth):
    """coors straight from generate(), full KITTI grid (1600 x 1408 map), the 70400-anchor car grid."""
    g = pkg.VoxelGenerator(KITTI["voxel_size"], KITTI["point_cloud_range"], 5, 20000)
    pts = synth.lidar_frame(60000, "kitti", seed=5)
    _, coors, _ = g.generate(torch.from_numpy(pts).cuda())
    bv = second_anchors_bv(KITTI["point_cloud_range"])
    assert bv.shape == (70400, 4)
    got = pkg.anchors_area(coors, bv, g.voxel_size, g.point_cloud_range, g.grid_size)
    want = oracle.anchors_area(coors.cpu().numpy(), bv, g.voxel_size, g.point_cloud_range, g.grid_size)
    assert np.array_equal(got.cpu().numpy(), want)
    assert 0 < int((want > 1).sum()) < want.size


def test_batched_matches_per_frame(pkg, oracle, synth):
    g = pkg.VoxelGenerator(KITTI["voxel_size"], KITTI["point_cloud_range"], 5, 6000)
    frames = [synth.lidar_frame(n, "kitti", seed=40 + i) for i, n in enumerate([20000, 0, 7000, 12000])]
    pts = torch.from_numpy(np.concatenate(frames)).cuda()
    offs = np.cumsum([0] + [f.shape[0] for f in frames]).astype(np.int32)
    _, coors4, _, voff = g.generate_batched(pts, offs)
    bv = second_anchors_bv(KITTI["point_cloud_range"], fm_xy=(88, 100), sizes=((1.6, 3.9), (0.6, 0.8)))
    got = pkg.anchors_area(coors4, bv, g.voxel_size, g.point_cloud_range, g.grid_size, batch_size=len(frames),
                           max_voxels_per_frame=6000).cpu().numpy()
    assert got.shape == (len(frames), bv.shape[0])
    c = coors4.cpu().numpy()
    voff = voff.cpu().numpy() if hasattr(voff, "cpu") else np.asarray(voff)
    for b in range(len(frames)):
        want = oracle.anchors_area(c[voff[b]:voff[b + 1], 1:], bv, g.voxel_size, g.point_cloud_range, g.grid_size)
        assert np.array_equal(got[b], want), b
    assert not got[1].any()                                   # the empty frame
    # same scratch again with fewer frames: the map must have been re-zeroed
    one = pkg.anchors_area(coors4[voff[2]:voff[3], 1:].contiguous(), bv, g.voxel_size, g.point_cloud_range,
                           g.grid_size).cpu().numpy()
    assert np.array_equal(one, got[2])


@pytest.mark.parametrize("case", ["odd_nx", "row_not_x4", "wide_cells", "tall", "wide_row", "tiny", "duplicates"])
def test_layouts_and_edges(pkg, oracle, case):
    """16-bit packed and 32-bit cells, rows that are / are not 16-byte multiples, many bands, a row so
    wide that a band is a single row, anchors sticking out of the range on every side."""
    rng = np.random.default_rng(sum(map(ord, case)))
    cap = None
    if case == "odd_nx":
        vs, rg = [0.5, 0.25, 1.0], [-10, -10, -2, 25.5, 10, 2]            # nx 71 -> 32-bit cells
    elif case == "row_not_x4":
        vs, rg = [0.5, 0.5, 1.0], [0, 0, 0, 19, 30, 2]                     # nx 38 -> 19 words per row
    elif case == "wide_cells":
        vs, rg = [0.25, 0.25, 0.25], [0, 0, 0, 40, 40, 10]                 # cap > 65535 -> 32-bit cells
        cap = 100000
    elif case == "tall":
        vs, rg = [1.0, 0.01, 1.0], [0, 0, 0, 24, 40, 1]                    # ny 4000: 125 bands of 32 rows
    elif case == "wide_row":
        vs, rg = [0.01, 1.0, 1.0], [0, 0, 0, 300.01, 3, 1]                 # nx 30001: 120 KB rows, 1-row bands
    elif case == "tiny":
        vs, rg = [1.0, 1.0, 1.0], [0, 0, 0, 2, 1, 1]                       # 1 x 2 map
    else:
        vs, rg = [0.5, 0.5, 0.1], [-4, -4, 0, 4, 4, 4]                     # 40 z levels over a 16 x 16 map
    vs, rg = np.asarray(vs, np.float32), np.asarray(rg, np.float32)
    grid = oracle.grid_size(vs, rg)
    m = {"tiny": 2, "duplicates": 9000, "wide_cells": 80000}.get(case, 4000)
    lin = rng.choice(int(grid.prod()), size=min(m, int(grid.prod())), replace=False)
    z, r = np.divmod(lin, int(grid[1] * grid[0]))
    y, x = np.divmod(r, int(grid[0]))
    coors = np.stack([z, y, x], 1).astype(np.int32)
    a = 3000
    lo = np.stack([rng.uniform(rg[0] - 2, rg[3] - 0.05, a), rng.uniform(rg[1] - 2, rg[4] - 0.05, a)], 1)
    hi = np.maximum(lo + rng.uniform(0.02, 6, (a, 2)), rg[:2] + 0.02)       # keep an overlap with the range
    bv = np.concatenate([lo, hi], 1).astype(np.float32)
    got = pkg.anchors_area(torch.from_numpy(coors).cuda(), bv, vs, rg, grid, max_voxels_per_frame=cap)
    want = oracle.anchors_area(coors, bv, vs, rg, grid)
    assert np.array_equal(got.cpu().numpy(), want), case
    assert case == "tiny" or want.max() > 0      # (a 1-row map has no rectangle above row 0)


def test_empty_and_errors(pkg):
    vs, rg, grid = [0.5, 0.5, 1.0], [0, 0, 0, 8, 8, 1], [16, 16, 1]
    bv = np.asarray([[0, 0, 4, 4], [1, 1, 2, 2]], np.float32)
    none = torch.zeros((0, 3), dtype=torch.int32, device="cuda")
    assert pkg.anchors_area(none, bv, vs, rg, grid).cpu().tolist() == [0.0, 0.0]
    some = torch.tensor([[0, 1, 1], [0, 3, 3]], dtype=torch.int32, device="cuda")
    assert pkg.anchors_area(some, np.zeros((0, 4), np.float32), vs, rg, grid).shape == (0,)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        pkg.anchors_area(some.cpu(), bv, vs, rg, grid)
    with pytest.raises(ValueError):
        pkg.anchors_area(torch.zeros((3, 4), dtype=torch.int32, device="cuda"), bv, vs, rg, grid)   # no batch_size
    with pytest.raises(ValueError):
        pkg.anchors_area(some, bv[:, :3], vs, rg, grid)
    with pytest.raises(ValueError):
        pkg.anchors_area(some, bv, [0.0, 0.5, 1.0], rg, grid)
    # rows outside the map or the batch are ignored, like the BEV scatter does
    bad = torch.tensor([[0, 1, 1], [0, 99, 3], [0, -1, 2]], dtype=torch.int32, device="cuda")
    assert pkg.anchors_area(bad, bv, vs, rg, grid).cpu().tolist() == [1.0, 0.0]
