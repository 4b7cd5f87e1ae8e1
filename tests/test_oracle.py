"""CPU tests of the oracle (the checker): three independent restatements of
SURVEY.md Appendix A must agree, the in-tree facts that constrain the voxelizer
 must hold, and the mean VFE / scatter restatements
must reproduce the golden vectors generated from the reference's own code."""
import os

import numpy as np
import pytest

COARSE = dict(voxel_size=[2.0, 2.0, 2.0], point_cloud_range=[-8, -8, -2, 8, 8, 2])
KITTI = dict(voxel_size=[0.05, 0.05, 0.1], point_cloud_range=[0, -40, -3, 70.4, 40, 1])


def _same(a, b):
    assert a[0].shape == b[0].shape
    assert np.array_equal(a[1], b[1]) and np.array_equal(a[2], b[2])
    assert np.array_equal(a[0].view(np.uint32), b[0].view(np.uint32))


def test_hand_checked_golden(oracle):
    """Tiny cloud worked by hand (grid 8x8x2 of 2 m cells, T=2, cap=3)."""
    pts = np.array([
        [0.5, 0.5, 0.5, 10],    # cell x=4 y=4 z=1 -> voxel 0
        [-7.5, -7.5, -1.5, 11],  # cell x=0 y=0 z=0 -> voxel 1
        [9.0, 0.0, 0.0, 12],    # outside (x >= 8)
        [1.9, 1.9, 1.9, 13],    # voxel 0, slot 1
        [0.1, 0.1, 0.1, 14],    # voxel 0 full (T=2) -> dropped
        [7.9, -8.0, 0.0, 15],   # cell x=7 y=0 z=1 -> voxel 2
        [2.0, 0.0, 0.0, 16],    # cell x=5 y=4 z=1 -> would be voxel 3 == cap -> overflow
        [-7.0, -7.0, -1.0, 17],  # voxel 1 again (after the overflow point)
    ], dtype=np.float32)
    for ov, n1 in (("break", 1), ("continue", 2)):
        g = oracle.OracleVoxelGenerator(COARSE["voxel_size"], COARSE["point_cloud_range"], 2, 3, ov)
        v, c, n = g.generate(pts)
        assert c.tolist() == [[1, 4, 4], [0, 0, 0], [1, 0, 7]]          # z, y, x
        assert n.tolist() == [2, n1, 1]
        assert v[0, :, 3].tolist() == [10, 13]
        assert v[1, :, 3].tolist() == ([11, 0] if ov == "break" else [11, 17])
        assert v[2, :, 3].tolist() == [15, 0]
        assert c.dtype == np.int32 and n.dtype == np.int32 and v.dtype == np.float32
    # default cap of the reference's builder (second/builder/voxel_builder.py:27)
    assert oracle.OracleVoxelGenerator([1, 1, 1], [0, 0, 0, 4, 4, 4], 3)._max_voxels == 20000


def test_grid_size_formula(oracle):
    # second/kittiviewer/viewer.py:468-469; every reference config -> 1408 x 1600 x 40
    assert oracle.grid_size(KITTI["voxel_size"], KITTI["point_cloud_range"]).tolist() == [1408, 1600, 40]
    assert oracle.grid_size([0.16, 0.16, 4], [0, -39.68, -3, 69.12, 39.68, 1]).tolist() == [432, 496, 1]
    assert oracle.grid_size([0.1, 0.1, 0.15], [-75.2, -75.2, -2, 75.2, 75.2, 4]).tolist() == [1504, 1504, 40]
    from oracle import pure
    for vs, rg in ([[0.05, 0.05, 0.1], KITTI["point_cloud_range"]], [[0.3, 0.7, 1.1], [-1, -2, -3, 4, 5, 6]],
                   [[0.4, 0.4, 0.4], [0, 0, 0, 1, 1.4, 1.8]]):
        assert oracle.grid_size(vs, rg).tolist() == pure.grid_size(vs, rg).tolist()
    assert oracle.grid_size(vs, rg).dtype == np.int64


@pytest.mark.parametrize("overflow", ["break", "continue"])
def test_three_restatements_agree(oracle, overflow):
    from oracle import numba_twin, pure
    rng = np.random.default_rng(5)
    for trial in range(6):
        n = int(rng.integers(1, 700))
        f = int(rng.integers(3, 7))
        pts = rng.uniform(-9, 9, size=(n, f)).astype(np.float32)
        pts[:, 2] = rng.uniform(-2.5, 2.5, size=n)
        T = int(rng.integers(1, 6))
        cap = int(rng.integers(1, 140))
        g = oracle.OracleVoxelGenerator(COARSE["voxel_size"], COARSE["point_cloud_range"], T, cap, overflow)
        a = g.generate(pts)
        b = pure.points_to_voxel(pts, COARSE["voxel_size"], COARSE["point_cloud_range"], T, cap, overflow)
        _same(a, b)
        if numba_twin.HAVE_NUMBA:
            c = numba_twin.points_to_voxel(pts, COARSE["voxel_size"], COARSE["point_cloud_range"], T, cap,
                                           overflow)
            _same(a, c)
        # the cached grid was restored to -1
        assert (g._grid == -1).all()


def test_kitti_shaped_c_vs_numba(oracle, synth):
    from oracle import numba_twin
    if not numba_twin.HAVE_NUMBA:
        pytest.skip("numba not importable")
    pts = synth.lidar_frame(40000, "kitti", seed=1)
    for ov in ("break", "continue"):
        g = oracle.OracleVoxelGenerator(KITTI["voxel_size"], KITTI["point_cloud_range"], 5, 6000, ov)
        _same(g.generate(pts), numba_twin.points_to_voxel(pts, KITTI["voxel_size"],
                                                          KITTI["point_cloud_range"], 5, 6000, ov))


def test_voxelizer_invariants(oracle, synth):
    """Facts fixed by the reference tree: zyx coords inside the grid,
    zero padding, counts in [1, T], first-come order, points inside their voxel."""
    pts = synth.lidar_frame(30000, "kitti", seed=2)
    g = oracle.OracleVoxelGenerator(KITTI["voxel_size"], KITTI["point_cloud_range"], 5, 20000)
    v, c, n = g.generate(pts)
    gs = g.grid_size
    assert (c >= 0).all() and (c < gs[::-1]).all()
    assert n.min() >= 1 and n.max() <= 5
    slot = np.arange(5)[None, :] < n[:, None]
    assert (v[~slot] == 0).all()
    vs, lo = g.voxel_size, g.point_cloud_range[:3]
    cell = np.floor((v[..., :3] - lo) / vs).astype(np.int32)[..., ::-1]
    assert (cell == c[:, None, :])[slot].all()
    # voxel_origins = coors[:, ::-1] * voxel_size + shift  (viewer.py:471-472)
    origin = c[:, ::-1] * vs + lo
    assert ((v[:, 0, :3] >= origin - 1e-4) & (v[:, 0, :3] <= origin + vs + 1e-4)).all()
    # first-come order: voxel k's first point appears before voxel k+1's first point
    key = {tuple(p): i for i, p in reversed(list(enumerate(map(tuple, pts))))}
    firsts = [key[tuple(v[k, 0])] for k in range(v.shape[0])]
    assert firsts == sorted(firsts)
    # unique voxels
    assert len({tuple(x) for x in c}) == c.shape[0]


def test_merge_sweeps_c_vs_numpy(oracle, synth):
    from oracle import pure
    sweeps = synth.lidar_sweeps(3, 2000, "nuscenes", 100, 0.05)
    a = oracle.merge_sweeps(sweeps)
    b = pure.merge_sweeps(sweeps)
    assert a.shape == (6000, 5)
    assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
    assert (a[:2000, 4] == 0).all() and np.allclose(a[4000:, 4], 0.1)
    assert np.array_equal(a[:2000, :4], sweeps[0][0])          # key frame copied verbatim


def test_mean_vfe_matches_reference_golden(oracle, golden_dir):
    from oracle import pure
    z = np.load(os.path.join(golden_dir, "vfe_reference.npz"))
    for tag in ("t5f4", "t100f4", "t10f5"):
        vox, num = z[f"{tag}_voxels"], z[f"{tag}_num"]
        got = oracle.mean_vfe(vox, num)
        assert got.shape == z[f"{tag}_v3"].shape          # [M, 4]: the reference's hard-coded slice
        np.testing.assert_allclose(got, z[f"{tag}_v3"], rtol=1e-6, atol=1e-6)
        np.testing.assert_allclose(pure.mean_vfe(vox, num), got, rtol=0, atol=0)


def test_scatter_matches_reference_golden(oracle, golden_dir):
    from oracle import pure
    z = np.load(os.path.join(golden_dir, "scatter_reference.npz"))
    for tag in ("pp_c4", "pp_c64", "d2_c3"):
        b, c, d, h, w = (int(x) for x in z[f"{tag}_shape"])
        got = oracle.scatter_bev(z[f"{tag}_feats"], z[f"{tag}_coors"], b, d, h, w)
        assert np.array_equal(got, z[f"{tag}_canvas"])
        assert np.array_equal(pure.scatter_bev(z[f"{tag}_feats"], z[f"{tag}_coors"], b, d, h, w), got)


def test_batched_oracle_contract(oracle, synth):
    frames = [synth.lidar_frame(n, "kitti", seed=i) for i, n in enumerate([3000, 0, 500])]
    g = oracle.OracleVoxelGenerator(KITTI["voxel_size"], KITTI["point_cloud_range"], 5, 400)
    v, c, n, off = g.generate_batched(frames)
    assert c.shape[1] == 4 and off.tolist()[0] == 0 and off[-1] == v.shape[0]
    assert off[1] == off[2]                                  # empty frame
    for b in range(3):
        assert (c[off[b]:off[b + 1], 0] == b).all()           # batch index first (voxelnet.py:285)


ANCHOR_TAGS = ("pillar", "second", "odd_nx")


def test_anchors_area_matches_reference_golden(oracle, golden_dir):
    """The restated sparse_sum_for_anchors_mask / fused_get_anchors_area against outputs of the
    reference's own functions (tests/golden/make_golden_anchors.py)."""
    z = np.load(os.path.join(golden_dir, "anchors_reference.npz"))
    for tag in ANCHOR_TAGS:
        grid, coors = z[f"{tag}_grid"], z[f"{tag}_coors"]
        hist = oracle.sparse_sum_for_anchors_mask(coors, (int(grid[1]), int(grid[0])))
        want = np.zeros_like(hist)
        nz = z[f"{tag}_hist_nonzero"]
        want[nz[:, 0], nz[:, 1]] = z[f"{tag}_hist_values"]
        assert np.array_equal(hist, want)
        area = oracle.anchors_area(coors, z[f"{tag}_anchors_bv"], z[f"{tag}_voxel_size"], z[f"{tag}_range"], grid)
        assert area.dtype == np.float32 and np.array_equal(area, z[f"{tag}_area"])
        # [M, 4] rows (b, z, y, x) read the same two columns
        c4 = np.concatenate([np.zeros((coors.shape[0], 1), np.int32), coors], 1)
        assert np.array_equal(oracle.anchors_area(c4, z[f"{tag}_anchors_bv"], z[f"{tag}_voxel_size"],
                                                  z[f"{tag}_range"], grid), z[f"{tag}_area"])


def test_anchors_area_is_a_box_count(oracle):
    """Size-independent property: the area is the number of voxels whose (y, x) lies in the
    half-open cell rectangle (c0, c2] x (c1, c3] -- the summed-area-table identity."""
    rng = np.random.default_rng(3)
    vs = np.asarray([0.25, 0.5, 1.0], np.float32)
    rg = np.asarray([-8, -6, -1, 8, 6, 1], np.float32)
    grid = oracle.grid_size(vs, rg)
    coors = np.stack([rng.integers(0, grid[2], 900), rng.integers(0, grid[1], 900),
                      rng.integers(0, grid[0], 900)], 1).astype(np.int32)
    lo = np.stack([rng.uniform(-8, 7, 300), rng.uniform(-6, 5, 300)], 1)
    bv = np.concatenate([lo, lo + rng.uniform(0.1, 4, (300, 2))], 1).astype(np.float32)
    got = oracle.anchors_area(coors, bv, vs, rg, grid)
    c = np.floor((bv - rg[[0, 1, 0, 1]]) / vs[[0, 1, 0, 1]]).astype(np.int64)
    c[:, 0] = np.maximum(c[:, 0], 0); c[:, 1] = np.maximum(c[:, 1], 0)
    c[:, 2] = np.minimum(c[:, 2], grid[0] - 1); c[:, 3] = np.minimum(c[:, 3], grid[1] - 1)
    y, x = coors[:, 1][None], coors[:, 2][None]
    want = ((x > c[:, 0:1]) & (x <= c[:, 2:3]) & (y > c[:, 1:2]) & (y <= c[:, 3:4])).sum(1)
    assert np.array_equal(got, want.astype(np.float32))


def test_crop_mask_matches_reference_golden(oracle, golden_dir):
    """The restated convex-polygon point test against outputs of the reference's own
    points_in_convex_polygon_3d_jit (tests/golden/make_golden_crop.py)."""
    z = np.load(os.path.join(golden_dir, "crop_reference.npz"))
    for tag in ("front", "left"):
        got = oracle.crop_mask(z[f"{tag}_points"], z[f"{tag}_normal_vec"], z[f"{tag}_d"])
        assert got.dtype == bool and np.array_equal(got, z[f"{tag}_mask"])
        assert 0 < got.sum() < got.size
