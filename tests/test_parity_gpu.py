"""GPU parity tests proper: the CUDA path (through the C ABI) against the CPU
oracle on the same seeded inputs.  Bar: bit-exact for voxels / coordinates /
counts / ordering (integer work and pure copies; the sweep transform has a fixed
operation order without FMA, so it is bit-exact too); 1e-6 relative for the
mean VFE and BEV canvases (fp32 sums)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")

KITTI = dict(voxel_size=[0.05, 0.05, 0.1], point_cloud_range=[0, -40, -3, 70.4, 40, 1])
PILLAR = dict(voxel_size=[0.16, 0.16, 4], point_cloud_range=[0, -39.68, -3, 69.12, 39.68, 1])
NUSC = dict(voxel_size=[0.1, 0.1, 0.2], point_cloud_range=[-51.2, -51.2, -5, 51.2, 51.2, 3])
COARSE = dict(voxel_size=[2.0, 2.0, 2.0], point_cloud_range=[-8, -8, -2, 8, 8, 2])


def _eq(got, want, what=""):
    got = [g.cpu().numpy() if hasattr(g, "cpu") else np.asarray(g) for g in got]
    assert got[0].shape == want[0].shape, f"{what}: M {got[0].shape} vs {want[0].shape}"
    assert np.array_equal(got[1], want[1]), f"{what}: coordinates differ"
    assert np.array_equal(got[2], want[2]), f"{what}: num_points differ"
    assert np.array_equal(got[0].view(np.uint32), want[0].view(np.uint32)), f"{what}: voxels differ"


def _pair(pkg, oracle, grid_cfg, T, cap, overflow="break", grid="hashed"):
    g = pkg.VoxelGenerator(grid_cfg["voxel_size"], grid_cfg["point_cloud_range"], T, cap,
                           overflow=overflow, grid=grid)
    o = oracle.OracleVoxelGenerator(grid_cfg["voxel_size"], grid_cfg["point_cloud_range"], T, cap,
                                    overflow=overflow)
    return g, o


@pytest.mark.parametrize("overflow", ["break", "continue"])
@pytest.mark.parametrize("grid", ["hashed", "dense"])
@pytest.mark.parametrize("n,cap", [(1, 10), (31, 10), (33, 1000), (1000, 50), (5000, 700), (20011, 3000),
                                   (20011, 100000)])
def test_kitti_shaped_device_form(pkg, oracle, synth, overflow, grid, n, cap):
    pts = synth.lidar_frame(n, "kitti", seed=n)
    g, o = _pair(pkg, oracle, KITTI, 5, cap, overflow, grid)
    want = o.generate(pts)
    got = g.generate(torch.from_numpy(pts).cuda())
    _eq(got, want, f"n={n} cap={cap} {overflow} {grid}")
    # same handle again: the workspace must have cleaned itself
    got = g.generate(torch.from_numpy(pts).cuda())
    _eq(got, want, "second call")


@pytest.mark.parametrize("overflow", ["break", "continue"])
def test_host_form_numpy_roundtrip(pkg, oracle, synth, overflow):
    pts = synth.lidar_frame(30000, "kitti", seed=3)
    g, o = _pair(pkg, oracle, KITTI, 5, 4000, overflow)
    got = g.generate(pts)
    assert isinstance(got[0], np.ndarray) and got[1].dtype == np.int32 and got[2].dtype == np.int32
    _eq(got, o.generate(pts), "host form")
    # call-time cap overrides the default like the reference's generate(points, max_voxels)
    _eq(g.generate(pts, 1234), o.generate(pts, 1234), "host form, call-time cap")
    # pinned input is used in place
    pp = pkg.pinned_empty(pts.shape, np.float32)
    pp[:] = pts
    _eq(g.generate(pp), o.generate(pts), "pinned host form")


@pytest.mark.parametrize("overflow", ["break", "continue"])
def test_many_more_cells_than_cap(pkg, oracle, synth, overflow):
    """Far more distinct cells than max_voxels (the table is sized by the point count, not by
    the cap); results must not change, and the next call must start from a clean workspace."""
    pts = synth.lidar_frame(40000, "kitti", seed=21)
    for cap in (64, 300, 5000):
        g, o = _pair(pkg, oracle, KITTI, 5, cap, overflow)
        want = o.generate(pts)
        for _ in range(2):
            _eq(g.generate(torch.from_numpy(pts).cuda()), want, f"cap={cap} {overflow}")
        small = synth.lidar_frame(900, "kitti", seed=22)          # same handle, tiny frame afterwards
        _eq(g.generate(torch.from_numpy(small).cuda()), o.generate(small), "after overflow")


def test_unshuffled_ring_order(pkg, oracle, synth):
    pts = synth.lidar_frame(40000, "kitti", seed=5, shuffle=False)
    for overflow in ("break", "continue"):
        g, o = _pair(pkg, oracle, KITTI, 5, 6000, overflow)
        _eq(g.generate(torch.from_numpy(pts).cuda()), o.generate(pts), f"ring order {overflow}")


@pytest.mark.parametrize("f", [3, 4, 5, 7])
def test_feature_counts(pkg, oracle, f):
    rng = np.random.default_rng(f)
    pts = rng.uniform(-9, 9, size=(7001, f)).astype(np.float32)
    g, o = _pair(pkg, oracle, COARSE, 6, 100)
    _eq(g.generate(torch.from_numpy(pts).cuda()), o.generate(pts), f"F={f}")
    _eq(g.generate(pts), o.generate(pts), f"F={f} host")


def test_empty_and_all_outside(pkg, oracle):
    g, o = _pair(pkg, oracle, COARSE, 4, 50)
    empty = np.zeros((0, 4), np.float32)
    got = g.generate(empty)
    assert got[0].shape == (0, 4, 4) and got[1].shape == (0, 3) and got[2].shape == (0,)
    got = g.generate(torch.from_numpy(empty).cuda())
    assert got[0].shape == (0, 4, 4)
    outside = np.full((100, 4), 1000.0, np.float32)
    _eq(g.generate(torch.from_numpy(outside).cuda()), o.generate(outside), "all outside")


def test_boundaries_nan_inf_negzero(pkg, oracle):
    vs, rg = COARSE["voxel_size"], COARSE["point_cloud_range"]
    rows = []
    for x in (-8.0, -8.000001, 7.999999, 8.0, -0.0, 0.0, 2.0, 1.9999999, np.nan, np.inf, -np.inf, 1e30):
        rows.append([x, 0.5, 0.5, 1.0])
        rows.append([0.5, x, 0.5, 2.0])
        rows.append([0.5, 0.5, x if abs(x) != 8.0 else x / 4, 3.0])
    pts = np.array(rows, dtype=np.float32)
    g, o = _pair(pkg, oracle, COARSE, 3, 50)
    _eq(g.generate(torch.from_numpy(pts).cuda()), o.generate(pts), "boundaries")
    # fp32 division corner: values on voxel edges of the KITTI grid
    g2, o2 = _pair(pkg, oracle, KITTI, 5, 20000)
    k = np.arange(0, 1408, dtype=np.float32)
    edge = np.stack([k * np.float32(0.05), np.zeros_like(k), np.zeros_like(k), k], 1)
    edge = np.concatenate([edge, np.nextafter(edge, np.float32(100)), np.nextafter(edge, np.float32(-100))])
    edge = np.ascontiguousarray(edge, dtype=np.float32)
    _eq(g2.generate(torch.from_numpy(edge).cuda()), o2.generate(edge), "voxel edges")


@pytest.mark.parametrize("overflow", ["break", "continue"])
def test_cap_exactly_and_one_over(pkg, oracle, overflow):
    # 12 distinct cells, several points each, interleaved
    rng = np.random.default_rng(0)
    cells = rng.permutation(np.repeat(np.arange(12), 7))
    pts = np.stack([cells % 4 * 2.0 - 7, cells // 4 * 2.0 - 7, np.zeros(len(cells)),
                    np.arange(len(cells))], 1).astype(np.float32)
    for cap in (11, 12, 13, 1):
        g, o = _pair(pkg, oracle, COARSE, 5, cap, overflow)
        _eq(g.generate(torch.from_numpy(pts).cuda()), o.generate(pts), f"cap={cap} {overflow}")


@pytest.mark.parametrize("T", [1, 2, 5, 8, 9, 35, 100])
def test_crowded_voxels_truncation(pkg, oracle, T):
    # everything in a handful of voxels: exercises max_points truncation in input order
    rng = np.random.default_rng(T)
    n = 30000
    pts = rng.uniform(-3.9, 3.9, size=(n, 4)).astype(np.float32)  # 16 cells of the coarse grid
    pts[:, 2] = 0.5
    g, o = _pair(pkg, oracle, COARSE, T, 40)
    _eq(g.generate(torch.from_numpy(pts).cuda()), o.generate(pts), f"T={T}")
    one = np.tile(np.array([[0.5, 0.5, 0.5, 0.0]], np.float32), (5000, 1))
    one[:, 3] = np.arange(5000)
    _eq(g.generate(torch.from_numpy(one).cuda()), o.generate(one), f"T={T} single voxel")
    _eq(g.generate(torch.from_numpy(pts).cuda()), o.generate(pts), f"T={T} repeat")


@pytest.mark.parametrize("overflow", ["break", "continue"])
def test_pillars_T100(pkg, oracle, synth, overflow):
    for shuffle in (True, False):
        pts = synth.lidar_frame(60000, "kitti", seed=11, shuffle=shuffle)
        g, o = _pair(pkg, oracle, PILLAR, 100, 5000, overflow)
        _eq(g.generate(torch.from_numpy(pts).cuda()), o.generate(pts), f"pillars shuffle={shuffle}")
        g, o = _pair(pkg, oracle, PILLAR, 12, 12000, overflow)   # many voxels over T
        _eq(g.generate(torch.from_numpy(pts).cuda()), o.generate(pts), f"pillars T=12 shuffle={shuffle}")


@pytest.mark.parametrize("overflow", ["break", "continue"])
def test_sweeps(pkg, oracle, synth, overflow):
    sweeps = synth.lidar_sweeps(4, 9000, "nuscenes", 100, 0.05)
    merged = oracle.merge_sweeps(sweeps)
    assert merged.shape == (36000, 5)
    g, o = _pair(pkg, oracle, NUSC, 10, 9000, overflow)
    want = o.generate(merged)
    dsw = [(torch.from_numpy(p).cuda(), R, t, dt) for p, R, t, dt in sweeps]
    _eq(g.generate_sweeps(dsw), want, "sweeps device")
    _eq(g.generate_sweeps(sweeps), want, "sweeps numpy")
    # ragged sweep sizes incl. an empty sweep, F_in = 5
    rng = np.random.default_rng(1)
    rag = []
    for s, n in enumerate([1000, 0, 37, 4001]):
        p = rng.uniform(-50, 50, size=(n, 5)).astype(np.float32)
        p[:, 2] = rng.uniform(-5, 3, size=n)
        R, t = synth.sweep_transform(s)
        rag.append((p, None if s == 0 else R, None if s == 0 else t, 0.05 * s))
    g, o = _pair(pkg, oracle, NUSC, 3, 2000, overflow)
    _eq(g.generate_sweeps(rag), o.generate(oracle.merge_sweeps(rag)), "ragged sweeps")


def _rot_z(a):
    c, s_ = np.cos(a), np.sin(a)
    return np.asarray([[c, -s_, 0], [s_, c, 0], [0, 0, 1]], np.float32)


@pytest.mark.parametrize("overflow", ["break", "continue"])
def test_batched_sweeps(pkg, oracle, synth, overflow):
    """B frames of 1-3 sweeps each (ragged, one empty sweep), dt appended: must equal the oracle run
    on the oracle's merged clouds, frame by frame, with the batch index in the coordinates."""
    rng = np.random.default_rng(8)
    frames = []
    for i, (k, n) in enumerate([(3, 9000), (1, 4000), (2, 0), (2, 12000)]):
        sw = []
        for j in range(k):
            pts = synth.lidar_frame(n if not (i == 3 and j == 1) else 0, "nuscenes", seed=900 + 10 * i + j)
            R = None if j == 0 else _rot_z(0.02 * j) @ np.eye(3, dtype=np.float32)
            t = None if j == 0 else rng.normal(0, 0.5, 3).astype(np.float32)
            sw.append((pts, R, t, 0.05 * j))
        frames.append(sw)
    for T, cap in ((5, 2500), (10, 60000)):
        g, o = _pair(pkg, oracle, NUSC, T, cap, overflow)
        merged = [oracle.merge_sweeps(sw) for sw in frames]
        wv, wc, wn, woff = o.generate_batched(merged)
        dev_frames = [[(torch.from_numpy(p).cuda(), R, t, dt) for p, R, t, dt in sw] for sw in frames]
        for _ in range(2):
            gv, gc, gn, goff = g.generate_batched_sweeps(dev_frames)
            assert np.array_equal(goff.cpu().numpy(), woff)
            _eq((gv, gc, gn), (wv, wc, wn), f"batched sweeps T={T}")
        # fused mean VFE over all five channels (dt included), padded form
        pv, pc, pn, poff, pmean = g.generate_batched_sweeps(dev_frames, mean_channels=5, padded=True)
        m = int(woff[-1])
        _eq((pv[:m], pc[:m], pn[:m]), (wv, wc, wn), f"batched sweeps T={T}, padded + mean")
        np.testing.assert_allclose(pmean[:m].cpu().numpy(), oracle.mean_vfe(wv, wn, 5), rtol=1e-6, atol=0)
        hv, hc, hn, hoff = g.generate_batched_sweeps(frames)              # numpy in -> numpy out
        assert isinstance(hv, np.ndarray) and np.array_equal(hoff, woff)
        _eq((hv, hc, hn), (wv, wc, wn), "batched sweeps, numpy")


@pytest.mark.parametrize("overflow", ["break", "continue"])
def test_shuffle_fused_into_the_load(pkg, oracle, synth, overflow):
    """The reference's shuffle_points (np.random.shuffle(points) before voxelization) as a permutation
    gathered on the point load: voxelizing (points, perm) must equal the oracle on points[perm] --
    plain frames, a frame that is also scaled, temporal sweeps with dt, T <= 8 and long buckets."""
    rng = np.random.default_rng(77)
    raw = [synth.lidar_frame(n, "kitti", seed=40 + i, shuffle=False) for i, n in enumerate([15000, 0, 7000, 1])]
    perms = [rng.permutation(p.shape[0]).astype(np.int32) for p in raw]
    S = np.eye(3, dtype=np.float32) * np.float32(0.95)
    xf = [(None, None), (None, None), (S, None), (None, None)]
    for T, cap in ((5, 1200), (20, 1200)):
        g, o = _pair(pkg, oracle, KITTI, T, cap, overflow)
        shuffled = [np.ascontiguousarray(p[q]) for p, q in zip(raw, perms)]
        merged = [np.ascontiguousarray(oracle.merge_sweeps([(p, R, t, 0.0)])[:, :4]) for p, (R, t) in zip(shuffled, xf)]
        want = o.generate_batched(merged)
        frames = [[(torch.from_numpy(p).cuda(), R, t, 0.0, torch.from_numpy(q).cuda())]
                  for p, q, (R, t) in zip(raw, perms, xf)]
        for _ in range(2):
            gv, gc, gn, goff = g.generate_batched_sweeps(frames, append_dt=False)
            assert np.array_equal(goff.cpu().numpy(), want[3])
            _eq((gv, gc, gn), want[:3], f"shuffle T={T}")
        hv, hc, hn, hoff = g.generate_batched_sweeps(
            [[(p, R, t, 0.0, q)] for p, q, (R, t) in zip(raw, perms, xf)], append_dt=False)   # numpy form
        _eq((hv, hc, hn), want[:3], f"shuffle numpy T={T}")
    # temporal: two shuffled sweeps with dt, single-frame form
    g, o = _pair(pkg, oracle, KITTI, 5, 2000, overflow)
    a, b = raw[0], raw[2]
    qa, qb = perms[0], perms[2]
    R, t = _rot_z(0.01), np.asarray([0.3, 0.0, 0.0], np.float32)
    wv, wc, wn = o.generate(oracle.merge_sweeps([(a[qa], None, None, 0.0), (b[qb], R, t, 0.1)]))
    gv, gc, gn = g.generate_sweeps([(torch.from_numpy(a).cuda(), None, None, 0.0, torch.from_numpy(qa).cuda()),
                                    (torch.from_numpy(b).cuda(), R, t, 0.1, torch.from_numpy(qb).cuda())])
    _eq((gv, gc, gn), (wv, wc, wn), "shuffled sweeps")
    with pytest.raises(ValueError):
        g.generate_sweeps([(a, None, None, 0.0, qa[:-1])])


def test_global_augmentation_fused_into_the_load(pkg, oracle, synth):
    """One transformed sweep per frame, no dt channel: the reference's global rotation / scaling / flip
    (second/core/preprocess.py:756-804) applied while the points are loaded.  Voxels are bit-exact
    against the oracle on the oracle's transformed points; the transformed coordinates themselves equal
    numpy's in-place float32 arithmetic exactly for scaling and flip, to 1e-6 for the rotation matmul."""
    g, o = _pair(pkg, oracle, KITTI, 5, 3000)
    pts = [synth.lidar_frame(15000, "kitti", seed=700 + i) for i in range(3)]
    scale = np.float32(1.0371)
    xf = [(np.eye(3, dtype=np.float32) * scale, None),                                   # global_scaling
          (np.diag([1, -1, 1]).astype(np.float32), None),                                # random_flip
          (_rot_z(0.3), None)]                                                           # global_rotation
    frames = [[(p, R, t, 0.0)] for p, (R, t) in zip(pts, xf)]
    merged = [np.ascontiguousarray(oracle.merge_sweeps(sw)[:, :4]) for sw in frames]
    want = o.generate_batched(merged)
    gv, gc, gn, goff = g.generate_batched_sweeps(
        [[(torch.from_numpy(p).cuda(), R, t, dt) for p, R, t, dt in sw] for sw in frames], append_dt=False)
    assert gv.shape[2] == 4 and np.array_equal(goff.cpu().numpy(), want[3])
    _eq((gv, gc, gn), want[:3], "augmentation fused into the load")
    ref0 = pts[0].copy(); ref0[:, :3] *= float(scale)                                    # preprocess.py:759
    ref1 = pts[1].copy(); ref1[:, 1] = -ref1[:, 1]                                       # preprocess.py:783
    ref2 = pts[2].copy(); ref2[:, :3] = ref2[:, :3] @ _rot_z(0.3).T                      # box_np_ops rotation, axis 2
    assert np.array_equal(np.abs(merged[0]), np.abs(ref0))       # (-0.0 vs 0.0 aside)
    assert np.array_equal(np.abs(merged[1]), np.abs(ref1))
    np.testing.assert_allclose(merged[2], ref2, rtol=1e-6, atol=1e-5)


@pytest.mark.parametrize("overflow", ["break", "continue"])
def test_frustum_crop_fused_into_the_load(pkg, oracle, synth, overflow, golden_dir):
    """remove_outside_points (second/core/box_np_ops.py:645-657) as crop planes riding the point load:
    dropping a point in the kernel must give the voxels the reference gets from the compacted array.
    Planes and the points sitting exactly on them come from the golden file the reference's own
    functions produced; one frame is also rotated after the crop, one is left uncropped."""
    import os
    z = np.load(os.path.join(golden_dir, "crop_reference.npz"))
    g, o = _pair(pkg, oracle, KITTI, 5, 900, overflow)
    frames, crops, want_frames = [], [], []
    for i, tag in enumerate(("front", "left", None)):
        if tag is None:
            pts, crop = synth.lidar_frame(7000, "kitti", seed=77), None
            keep = np.ones(pts.shape[0], bool)
        else:
            pts, crop = z[f"{tag}_points"], (z[f"{tag}_normal_vec"], z[f"{tag}_d"])
            keep = oracle.crop_mask(pts, *crop)
            assert np.array_equal(keep, z[f"{tag}_mask"])
        R = _rot_z(0.2) if i == 1 else None
        t = np.asarray([0.5, -0.25, 0.0], np.float32) if i == 1 else None
        frames.append([(pts, R, t, 0.0)])
        crops.append(crop)
        kept = np.ascontiguousarray(pts[keep])                       # what the reference voxelizes
        want_frames.append(np.ascontiguousarray(oracle.merge_sweeps([(kept, R, t, 0.0)])[:, :4]))
    want = o.generate_batched(want_frames)
    dev = [[(torch.from_numpy(np.ascontiguousarray(p)).cuda(), R, t, dt) for p, R, t, dt in sw] for sw in frames]
    for _ in range(2):
        gv, gc, gn, goff = g.generate_batched_sweeps(dev, append_dt=False, crops=crops)
        assert np.array_equal(goff.cpu().numpy(), want[3])
        _eq((gv, gc, gn), want[:3], "frustum crop")
    with pytest.raises(ValueError):
        g.generate_batched_sweeps(dev, append_dt=False, crops=crops[:2])
    with pytest.raises(ValueError):
        g.generate_batched_sweeps(dev, append_dt=False, crops=[(np.zeros((9, 3)), np.zeros(9)), None, None])


@pytest.mark.parametrize("overflow", ["break", "continue"])
def test_batched(pkg, oracle, synth, overflow):
    sizes = [5000, 0, 33, 12001, 1, 7000]
    frames = [synth.lidar_frame(n, "kitti", seed=40 + i) for i, n in enumerate(sizes)]
    g, o = _pair(pkg, oracle, KITTI, 5, 2500, overflow)
    wv, wc, wn, woff = o.generate_batched(frames)
    dv, dc, dn, doff = g.generate_batched([torch.from_numpy(f).cuda() for f in frames])
    assert np.array_equal(doff.cpu().numpy(), woff)
    _eq((dv, dc, dn), (wv, wc, wn), "batched device")
    hv, hc, hn, hoff = g.generate_batched(frames)
    assert np.array_equal(hoff, woff)
    _eq((hv, hc, hn), (wv, wc, wn), "batched host")
    # fused mean VFE == separate module == oracle
    out = g.generate_batched([torch.from_numpy(f).cuda() for f in frames], mean_channels=4)
    mean = out[4].cpu().numpy()
    want = oracle.mean_vfe(wv, wn)
    np.testing.assert_allclose(mean, want, rtol=1e-6, atol=0)
    sep = pkg.VoxelFeatureExtractorV3()(dv, dn).cpu().numpy()
    assert np.array_equal(sep.view(np.uint32), mean.view(np.uint32))


@pytest.mark.parametrize("overflow", ["break", "continue"])
def test_batched_host_pipelined(pkg, oracle, synth, overflow):
    """Pinned host buffers take the pipelined path (sub-batches on three streams); results
    must be identical to the plain host form, for ragged batches with empty frames too."""
    for sizes in ([5000, 0, 33, 12001, 1, 7000, 0, 0, 2500, 900, 15000], [300, 200], [0, 0, 0],
                  [4000] * 19):
        frames = [synth.lidar_frame(n, "kitti", seed=70 + i) for i, n in enumerate(sizes)]
        g, o = _pair(pkg, oracle, KITTI, 5, 1500, overflow)
        wv, wc, wn, woff = o.generate_batched(frames)
        cat = np.concatenate(frames, axis=0)
        offs = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
        hp = pkg.pinned_empty(cat.shape, np.float32)
        hp[:] = cat
        b, cap = len(sizes), 1500
        out = (pkg.pinned_empty((b * cap, 5, 4), np.float32), pkg.pinned_empty((b * cap, 4), np.int32),
               pkg.pinned_empty((b * cap,), np.int32), pkg.pinned_empty((b + 1,), np.int32))
        for _ in range(2):                      # second call: workspace and events reused
            hv, hc, hn, hoff = g.generate_batched(hp, offs, out=out)
            assert np.array_equal(hoff, woff), sizes
            _eq((hv, hc, hn), (wv, wc, wn), f"pipelined host {sizes[:3]}")


@pytest.mark.parametrize("T,f", [(5, 4), (3, 5), (35, 4), (12, 7)])
@pytest.mark.parametrize("assemble", [True, False])
def test_host_assembly_matches_device_rows(pkg, oracle, synth, T, f, assemble):
    """Pinned batched host form with host assembly (point indices come home, the padded rows are
    copied from the caller's points by host threads) against the oracle, for the specialised
    gather (T <= 8, 4 features) and the general one (long buckets, other feature counts), ragged
    batches with empty frames, crowded voxels; and the same with the device-side rows."""
    rng = np.random.default_rng(T * 100 + f)
    cfg = dict(voxel_size=[0.4, 0.4, 0.8], point_cloud_range=[0, -40, -3, 70.4, 40, 1])
    cap = 900
    sizes = [7000, 0, 1, 9000, 350, 0, 12000, 5000, 2, 3000, 8000]
    frames = []
    for i, n in enumerate(sizes):
        p4 = synth.lidar_frame(n, "kitti", seed=300 + i)
        extra = rng.normal(0, 1, size=(n, f - 4)).astype(np.float32) if f > 4 else np.zeros((n, 0), np.float32)
        frames.append(np.ascontiguousarray(np.concatenate([p4, extra], axis=1)[:, :f]))
    for overflow in ("break", "continue"):
        g = pkg.VoxelGenerator(cfg["voxel_size"], cfg["point_cloud_range"], T, cap, overflow=overflow,
                               host_assemble=assemble)
        o = oracle.OracleVoxelGenerator(cfg["voxel_size"], cfg["point_cloud_range"], T, cap, overflow)
        wv, wc, wn, woff = o.generate_batched(frames)
        cat = np.concatenate(frames, axis=0)
        offs = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
        hp = pkg.pinned_empty(cat.shape, np.float32)
        hp[:] = cat
        b = len(sizes)
        out = (pkg.pinned_empty((b * cap, T, f), np.float32), pkg.pinned_empty((b * cap, 4), np.int32),
               pkg.pinned_empty((b * cap,), np.int32), pkg.pinned_empty((b + 1,), np.int32))
        for rep in range(3):
            out[0][:] = np.float32(7.0)                 # stale rows must be overwritten, padding included
            hv, hc, hn, hoff = g.generate_batched(hp, offs, out=out)
            assert np.array_equal(hoff, woff)
            _eq((hv, hc, hn), (wv, wc, wn), f"host assembly={assemble} T={T} f={f} {overflow} rep {rep}")
        h2d, d2h, _ = g.last_transfer()
        m = int(woff[-1])
        if not assemble:
            bound = m * (T * 4 * f + 20)
        elif T <= 8 and f == 4:              # compact kept rows + one int per gather tile
            bound = int(wn.sum()) * 16 + m * 20 + b * ((cap + 255) // 256) * 4
        else:                                # point indices
            bound = m * (T * 4 + 20)
        assert d2h <= bound + 4096, (d2h, bound)
        g.close()


def _pinned_batch(pkg, frames, cap, T=5):
    cat = np.concatenate(frames, axis=0)
    offs = np.concatenate([[0], np.cumsum([f.shape[0] for f in frames])]).astype(np.int32)
    hp = pkg.pinned_empty(cat.shape, np.float32)
    hp[:] = cat
    b = len(frames)
    out = (pkg.pinned_empty((b * cap, T, 4), np.float32), pkg.pinned_empty((b * cap, 4), np.int32),
           pkg.pinned_empty((b * cap,), np.int32), pkg.pinned_empty((b + 1,), np.int32))
    return hp, offs, out


def test_prefix_upload_is_exact(pkg, oracle, synth):
    """Break mode, pinned batches: after the first call only a prefix of each frame is uploaded
    (sized from the previous call's cut-offs).  Results must equal the oracle's whether the guess
    holds (later calls on similar frames), fails for every frame (a call whose cap is never
    reached) or fails for some frames of a sub-batch (mixed batch)."""
    cap = 1500
    g, o = _pair(pkg, oracle, KITTI, 5, cap)
    early = [synth.lidar_frame(30000, "kitti", seed=300 + i) for i in range(12)]      # cut-off near 10 %
    hp, offs, out = _pinned_batch(pkg, early, cap)
    want = o.generate_batched(early)
    full = hp.nbytes
    seen = []
    for _ in range(3):
        hv, hc, hn, hoff = g.generate_batched(hp, offs, out=out)
        assert np.array_equal(hoff, want[3])
        _eq((hv, hc, hn), want[:3], "prefix upload, guess holds")
        seen.append(g.last_transfer())
    assert seen[0][0] == full and seen[0][2] == 0                 # first call: everything
    assert seen[1][0] < 0.5 * full and seen[1][2] == 0            # then a prefix, nothing redone
    assert seen[2][0] == seen[1][0]
    # same handle, frames that never reach the cap: every prefix is too short, all sub-batches redone
    sparse = [synth.lidar_frame(30000, "kitti", seed=400 + i)[::40].copy() for i in range(12)]
    sparse = [np.ascontiguousarray(np.repeat(f, 40, axis=0)[:30000]) for f in sparse]  # 750 distinct points each
    hp2, offs2, out2 = _pinned_batch(pkg, sparse, cap)
    want2 = o.generate_batched(sparse)
    assert want2[3][-1] < 12 * cap                                  # (no frame is cut)
    hv, hc, hn, hoff = g.generate_batched(hp2, offs2, out=out2)
    assert np.array_equal(hoff, want2[3])
    _eq((hv, hc, hn), want2[:3], "prefix upload, guess fails everywhere")
    h2d, _, redone = g.last_transfer()
    assert redone > 0 and h2d == hp2.nbytes                         # the rest of each frame followed
    # the guess is now 'no cut-off': the next call uploads everything again and redoes nothing
    hv, hc, hn, hoff = g.generate_batched(hp, offs, out=out)
    _eq((hv, hc, hn), want[:3], "back to cut frames")
    assert g.last_transfer()[0] == full and g.last_transfer()[2] == 0
    g.generate_batched(hp, offs, out=out)                           # learn the early cut-off again
    # mixed batch: cut and uncut frames side by side in one sub-batch
    mixed = [early[0], sparse[0], early[1], sparse[1], sparse[2], early[2], early[3], sparse[3], early[4]]
    hp3, offs3, out3 = _pinned_batch(pkg, mixed, cap)
    want3 = o.generate_batched(mixed)
    hv, hc, hn, hoff = g.generate_batched(hp3, offs3, out=out3)
    assert np.array_equal(hoff, want3[3])
    _eq((hv, hc, hn), want3[:3], "prefix upload, mixed batch")
    assert g.last_transfer()[2] > 0
    # switched off: always the whole frame
    g2 = pkg.VoxelGenerator(KITTI["voxel_size"], KITTI["point_cloud_range"], 5, cap, prefix_upload=False)
    for _ in range(2):
        hv, hc, hn, hoff = g2.generate_batched(hp, offs, out=out)
        assert g2.last_transfer()[0] == full
    _eq((hv, hc, hn), want[:3], "prefix upload off")


def test_collate_points_feeds_the_host_form(pkg, oracle, synth):
    frames = [synth.lidar_frame(n, "kitti", seed=600 + i) for i, n in enumerate([4000, 0, 2500, 9000])]
    pts, offs = pkg.collate_points(frames)
    assert pts.shape == (15500, 4) and offs.tolist() == [0, 4000, 4000, 6500, 15500]
    g, o = _pair(pkg, oracle, KITTI, 5, 1200)
    want = o.generate_batched(frames)
    got = g.generate_batched(pts, offs)
    assert np.array_equal(got[3], want[3])
    _eq(got[:3], want[:3], "collate_points -> generate_batched")
    with pytest.raises(ValueError):
        pkg.collate_points([frames[0], frames[0][:, :3]])


def test_pageable_after_pinned_keeps_streams(pkg, oracle, synth):
    """Growing the pageable staging buffer must not disturb the pipelined form's streams and events."""
    cap = 1500
    g, o = _pair(pkg, oracle, KITTI, 5, cap)
    frames = [synth.lidar_frame(6000, "kitti", seed=500 + i) for i in range(5)]
    hp, offs, out = _pinned_batch(pkg, frames, cap)
    want = o.generate_batched(frames)
    _eq(g.generate_batched(hp, offs, out=out)[:3], want[:3], "pinned first")
    same = np.concatenate(frames, axis=0)                           # pageable, same handle: staging buffer is created
    _eq(g.generate_batched(same, offs)[:3], want[:3], "pageable, same handle")
    _eq(g.generate_batched(hp, offs, out=out)[:3], want[:3], "pinned after pageable")
    big = np.concatenate(frames + frames, axis=0)                   # pageable, larger than anything staged so far
    offs_big = np.concatenate([[0], np.cumsum([6000] * 10)]).astype(np.int32)
    wb = o.generate_batched(frames + frames)
    _eq(g.generate_batched(big, offs_big)[:3], wb[:3], "pageable, staging buffer grows")
    hp, offs, out = _pinned_batch(pkg, frames, cap)                 # (the handle may have been re-created: fresh buffers)
    _eq(g.generate_batched(hp, offs, out=out)[:3], want[:3], "pinned again")
    _eq(g.generate_batched(hp, offs, out=out)[:3], want[:3], "pinned again, second call")


@pytest.mark.parametrize("seed", range(6))
def test_randomized_configurations(pkg, oracle, synth, seed):
    """Seeded sweep over grids, caps, T, feature counts, batch shapes and both overflow rules, several
    calls per handle (everything the pipeline empties for itself must really be empty again)."""
    rng = np.random.default_rng(1000 + seed)
    for _ in range(5):
        vs = rng.choice([0.05, 0.1, 0.16, 0.25, 0.5, 1.0], 3).astype(np.float32)
        vs[2] = rng.choice([0.1, 0.2, 0.5, 4.0])
        cells = rng.integers(4, 400, 3)
        lo = rng.uniform(-60, 0, 3).astype(np.float32)
        rg = np.concatenate([lo, lo + vs * cells]).astype(np.float32)
        T = int(rng.choice([1, 2, 3, 5, 8, 9, 12, 35]))
        cap = int(rng.choice([1, 7, 100, 1000, 5000, 40000]))
        f = int(rng.choice([3, 4, 4, 4, 5]))
        overflow = str(rng.choice(["break", "continue"]))
        g = pkg.VoxelGenerator(vs, rg, T, cap, overflow=overflow)
        o = oracle.OracleVoxelGenerator(vs, rg, T, cap, overflow=overflow)
        for _call in range(3):
            b = int(rng.integers(1, 10))
            frames = []
            for _i in range(b):
                n = int(rng.choice([0, 1, 31, 1024, 1025, 3000, 20000, 45000]))
                pts = rng.uniform(rg[:3] - 2 * vs, rg[3:] + 2 * vs, (n, 3)).astype(np.float32)
                if n and rng.random() < 0.5:                  # crowd a few cells
                    pts[: n // 2] = pts[rng.integers(0, max(n // 50, 1), n // 2)] + \
                        rng.uniform(0, 0.2, (n // 2, 3)).astype(np.float32) * vs
                extra = rng.uniform(0, 1, (n, f - 3)).astype(np.float32)
                frames.append(np.ascontiguousarray(np.concatenate([pts, extra], 1)))
            want = o.generate_batched(frames)
            cat = torch.from_numpy(np.concatenate(frames, 0)).cuda()
            offs = np.concatenate([[0], np.cumsum([x.shape[0] for x in frames])]).astype(np.int32)
            gv, gc, gn, goff = g.generate_batched(cat, offs)[:4]
            what = f"seed={seed} vs={vs} T={T} cap={cap} f={f} {overflow} sizes={[x.shape[0] for x in frames]}"
            assert np.array_equal(goff.cpu().numpy(), want[3]), what
            _eq((gv, gc, gn), want[:3], what)
            if b == 1:
                _eq(g.generate(cat), o.generate(frames[0]), what + " single")


def test_handle_growth_and_reuse(pkg, oracle, synth):
    g, o = _pair(pkg, oracle, KITTI, 5, 2000)
    for n in (1000, 50000, 200, 80000, 50000):
        pts = synth.lidar_frame(n, "kitti", seed=n)
        _eq(g.generate(torch.from_numpy(pts).cuda()), o.generate(pts), f"growth n={n}")
    _eq(g.generate(pts, 9000), o.generate(pts, 9000), "bigger cap")


@pytest.mark.parametrize("name", ["cfg1", "cfg3"])
@pytest.mark.parametrize("overflow", ["break", "continue"])
def test_full_size_single_frame_configs(pkg, oracle, synth, name, overflow):
    c = synth.CONFIGS[name]
    pts = synth.make_frames(name, 1)[0]
    g = pkg.VoxelGenerator(c["voxel_size"], c["point_cloud_range"], c["max_num_points"], c["max_voxels"],
                           overflow=overflow)
    o = oracle.OracleVoxelGenerator(c["voxel_size"], c["point_cloud_range"], c["max_num_points"],
                                    c["max_voxels"], overflow=overflow)
    _eq(g.generate(torch.from_numpy(pts).cuda()), o.generate(pts), name)


@pytest.mark.parametrize("name", ["cfg2", "cfg5"])
@pytest.mark.parametrize("overflow", ["break", "continue"])
def test_full_size_sweep_configs(pkg, oracle, synth, name, overflow):
    """cfg2 (5 sweeps, T=10) and cfg5 (3 x 180k points, 150k-voxel cap: with `continue` every one of the
    540k points goes through the hashed table) at full size, both overflow rules."""
    c = synth.CONFIGS[name]
    sweeps = synth.make_sweeps(name)
    g = pkg.VoxelGenerator(c["voxel_size"], c["point_cloud_range"], c["max_num_points"], c["max_voxels"],
                           overflow=overflow)
    o = oracle.OracleVoxelGenerator(c["voxel_size"], c["point_cloud_range"], c["max_num_points"],
                                    c["max_voxels"], overflow=overflow)
    dsw = [(torch.from_numpy(p).cuda(), R, t, dt) for p, R, t, dt in sweeps]
    want = o.generate(oracle.merge_sweeps(sweeps))
    _eq(g.generate_sweeps(dsw), want, name)
    _eq(g.generate_sweeps(dsw), want, name + " second call")


def _cfg4_oracle(oracle, synth, overflow):
    c = synth.CONFIGS["cfg4"]
    frames = synth.make_frames("cfg4", 64)
    o = oracle.OracleVoxelGenerator(c["voxel_size"], c["point_cloud_range"], c["max_num_points"],
                                    c["max_voxels"], overflow=overflow)
    return c, frames, o.generate_batched(frames)


_CFG4_CACHE = {}


def _cfg4(oracle, synth, overflow):
    if overflow not in _CFG4_CACHE:
        _CFG4_CACHE[overflow] = _cfg4_oracle(oracle, synth, overflow)
    return _CFG4_CACHE[overflow]


@pytest.mark.parametrize("overflow", ["break", "continue"])
def test_full_size_cfg4_device_form(pkg, oracle, synth, overflow):
    """BASELINE cfg4 (64 x 120k points), the batch bench.py times on the device: ALL 64 frames against the
    oracle, both overflow rules, fused mean VFE included, twice on one handle; plus size-independent
    properties (sorted offsets, counts in [1, T], every kept point inside its voxel, zero padding)."""
    c, frames, want = _cfg4(oracle, synth, overflow)
    g = pkg.VoxelGenerator(c["voxel_size"], c["point_cloud_range"], c["max_num_points"], c["max_voxels"],
                           overflow=overflow)
    cat = torch.from_numpy(np.concatenate(frames)).cuda()
    offs = np.concatenate([[0], np.cumsum([f.shape[0] for f in frames])]).astype(np.int32)
    for call in range(2):
        v, cc, n, off, mean = g.generate_batched(cat, offs, mean_channels=4)
        assert np.array_equal(off.cpu().numpy(), want[3]), f"offsets, call {call}"
        _eq((v, cc, n), want[:3], f"cfg4 device form {overflow}, call {call}")
        np.testing.assert_allclose(mean.cpu().numpy(), oracle.mean_vfe(want[0], want[2]), rtol=1e-6, atol=0)
    off = want[3]
    assert off[0] == 0 and np.all(np.diff(off) == c["max_voxels"])   # the cap is hit in every frame
    vn, cn, nn = v.cpu().numpy(), cc.cpu().numpy(), n.cpu().numpy()
    assert nn.min() >= 1 and nn.max() <= c["max_num_points"]
    assert np.array_equal(cn[:, 0], np.repeat(np.arange(64), c["max_voxels"]))
    vs = np.asarray(c["voxel_size"], np.float32)
    lo = np.asarray(c["point_cloud_range"][:3], np.float32)
    slot = np.arange(c["max_num_points"])[None, :] < nn[:, None]
    cell = np.floor((vn[..., :3] - lo) / vs).astype(np.int32)[..., ::-1]
    assert np.all((cell == cn[:, None, 1:])[slot])
    assert np.all(vn[~slot] == 0)


@pytest.mark.parametrize("overflow", ["break", "continue"])
@pytest.mark.parametrize("prefix_upload,host_assemble", [(True, True), (True, False), (False, True), (False, False)])
def test_full_size_cfg4_host_form(pkg, oracle, synth, overflow, prefix_upload, host_assemble):
    """The path bench.py's end-to-end figure runs -- generate_batched(pinned numpy) at 64 x 120k points:
    prefix upload x host assembly on and off, both overflow rules, all 64 frames against the oracle.  Three
    calls on one handle: the first uploads whole frames, the second works from the guessed prefix, the
    third gets the frames in another order (its guess comes from different frames)."""
    c, frames, want = _cfg4(oracle, synth, overflow)
    cap, T = c["max_voxels"], c["max_num_points"]
    g = pkg.VoxelGenerator(c["voxel_size"], c["point_cloud_range"], T, cap, overflow=overflow,
                           prefix_upload=prefix_upload, host_assemble=host_assemble)
    hp, offs, out = _pinned_batch(pkg, frames, cap, T)
    for call in range(2):
        hv, hc, hn, hoff = g.generate_batched(hp, offs, out=out)
        assert np.array_equal(hoff, want[3])
        _eq((hv, hc, hn), want[:3], f"cfg4 host form {overflow} prefix={prefix_upload} assemble={host_assemble} call {call}")
        h2d, d2h, redone = g.last_transfer()
        if call == 1 and overflow == "break" and prefix_upload:
            assert h2d < 0.8 * hp.nbytes and redone == 0          # the guess held: a prefix was enough
        if call == 0 or overflow == "continue" or not prefix_upload:
            assert h2d == hp.nbytes
    order = np.roll(np.arange(64), -17)
    hp2, offs2, out2 = _pinned_batch(pkg, [frames[i] for i in order], cap, T)
    hv, hc, hn, hoff = g.generate_batched(hp2, offs2, out=out2)
    o = want[3]
    for k, i in enumerate(order):                                   # frame k of this call is frame i of the oracle's
        a, b = hoff[k], hoff[k + 1]
        assert b - a == o[i + 1] - o[i]
        assert np.all(hc[a:b, 0] == k)
        _eq((hv[a:b], hc[a:b, 1:], hn[a:b]), (want[0][o[i]:o[i + 1]], want[1][o[i]:o[i + 1], 1:], want[2][o[i]:o[i + 1]]),
            f"rotated batch, frame {k}")


def test_table_load_factor_worst_case(pkg, oracle):
    """N = 2^k - 1 points, every one in a cell of its own, continue mode with a cap above N: the hashed table is
    sized so that its load factor stays below 11/12 (it was N / 2^k ~ 1.0, i.e. O(N) probes per point)."""
    n = (1 << 14) - 1
    grid = dict(voxel_size=[1.0, 1.0, 1.0], point_cloud_range=[0, 0, 0, 128, 128, 1])
    ij = np.arange(n)
    pts = np.stack([ij % 128 + 0.5, ij // 128 + 0.5, np.full(n, 0.5), np.zeros(n)], axis=1).astype(np.float32)
    pts = pts[np.random.default_rng(0).permutation(n)]
    for overflow in ("continue", "break"):
        g, o = _pair(pkg, oracle, grid, 3, 20000, overflow)
        _eq(g.generate(torch.from_numpy(pts).cuda()), o.generate(pts), f"all distinct, {overflow}")


def test_big_frame_scan_path(pkg, oracle, synth, monkeypatch):
    """Frames of more than 4096 chunks (> 4 M points) scan their chunk counts in a kernel of their own
    (k_scan_big) instead of in shared memory; the limit is lowered through BVX_SCAN_SMEM_CHUNKS so that the
    path runs on a small frame: same results."""
    monkeypatch.setenv("BVX_SCAN_SMEM_CHUNKS", "4")
    frames = [synth.lidar_frame(n, "kitti", seed=900 + i) for i, n in enumerate([30000, 5000, 0, 21000])]
    for overflow in ("break", "continue"):
        for T, cap in ((5, 1500), (12, 900), (1, 100000)):
            g, o = _pair(pkg, oracle, KITTI, T, cap, overflow)
            want = o.generate_batched(frames)
            for _ in range(2):
                got = g.generate_batched([torch.from_numpy(f).cuda() for f in frames])
                assert np.array_equal(got[3].cpu().numpy(), want[3])
                _eq(got[:3], want[:3], f"big-frame path T={T} cap={cap} {overflow}")
            g.close()


def test_mean_vfe_and_scatter_golden(pkg, oracle, golden_dir):
    import os
    z = np.load(os.path.join(golden_dir, "vfe_reference.npz"))
    for tag in ("t5f4", "t100f4", "t10f5"):
        vox, num = z[f"{tag}_voxels"], z[f"{tag}_num"]
        got = pkg.VoxelFeatureExtractorV3()(torch.from_numpy(vox).cuda(), torch.from_numpy(num).cuda())
        np.testing.assert_allclose(got.cpu().numpy(), z[f"{tag}_v3"], rtol=1e-6, atol=1e-6)
        got = pkg.SimpleVoxel()(torch.from_numpy(vox).cuda(), torch.from_numpy(num).cuda())
        np.testing.assert_allclose(got.cpu().numpy(), z[f"{tag}_simple"], rtol=1e-6, atol=1e-6)
        # all channels (keeps dt for temporal input) vs oracle
        f = vox.shape[2]
        got = pkg.VoxelFeatureExtractorV3(num_output_channels=f)(torch.from_numpy(vox).cuda(),
                                                                 torch.from_numpy(num).cuda())
        np.testing.assert_allclose(got.cpu().numpy(), oracle.mean_vfe(vox, num, f), rtol=1e-6, atol=0)
    z = np.load(os.path.join(golden_dir, "scatter_reference.npz"))
    for tag in ("pp_c4", "pp_c64", "d2_c3"):
        b, c, d, h, w = (int(x) for x in z[f"{tag}_shape"])
        mod = pkg.PointPillarsScatter([1, d, h, w, c], num_input_features=c)
        got = mod(torch.from_numpy(z[f"{tag}_feats"]).cuda(), torch.from_numpy(z[f"{tag}_coors"]).cuda(), b)
        assert got.shape == (b, c * d, h, w)
        assert np.array_equal(got.cpu().numpy(), z[f"{tag}_canvas"])


def test_pillar_pipeline_to_canvas(pkg, oracle, synth):
    c = synth.CONFIGS["cfg3"]
    frames = [synth.lidar_frame(50000, "kitti", seed=90 + i) for i in range(3)]
    g = pkg.VoxelGenerator(c["voxel_size"], c["point_cloud_range"], c["max_num_points"], c["max_voxels"])
    o = oracle.OracleVoxelGenerator(c["voxel_size"], c["point_cloud_range"], c["max_num_points"],
                                    c["max_voxels"])
    v, cc, n, off = g.generate_batched([torch.from_numpy(f).cuda() for f in frames])
    wv, wc, wn, woff = o.generate_batched(frames)
    _eq((v, cc, n), (wv, wc, wn), "pillars batched")
    feat = pkg.VoxelFeatureExtractorV3()(v, n)
    gx, gy, gz = (int(x) for x in g.grid_size)
    canvas = pkg.PointPillarsScatter([1, gz, gy, gx, 4])(feat, cc, 3)
    want = oracle.scatter_bev(oracle.mean_vfe(wv, wn), wc, 3, gz, gy, gx)
    np.testing.assert_allclose(canvas.cpu().numpy(), want, rtol=1e-6, atol=0)
    # backward of the scatter is the gather of the written cells
    feat2 = feat.clone().requires_grad_(True)
    out = pkg.PointPillarsScatter([1, gz, gy, gx, 4])(feat2, cc, 3)
    out.sum().backward()
    assert torch.all(feat2.grad == 1)


@pytest.mark.parametrize("name,T,cap", [("cfg3", 100, 12000), ("cfg3", 100, 700), ("coarse", 5, 3000)])
def test_fused_points_to_canvas(pkg, oracle, synth, name, T, cap):
    """generate_bev: voxelization + fused mean VFE + BEV scatter with the voxel count read on the device,
    the padded rows never written.  Must equal the three separate reference steps on the oracle's voxels
    (pillars through the general gather, SECOND-sized voxels through the specialised one, D > 1)."""
    c = synth.CONFIGS[name] if name != "coarse" else dict(voxel_size=[0.4, 0.4, 1.0],
                                                         point_cloud_range=[0, -40, -3, 70.4, 40, 1])
    sizes = [50000, 0, 30000, 1]
    frames = [synth.lidar_frame(n, "kitti", seed=120 + i) for i, n in enumerate(sizes)]
    offs = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
    cat = torch.from_numpy(np.concatenate(frames)).cuda()
    for overflow in ("break", "continue"):
        g = pkg.VoxelGenerator(c["voxel_size"], c["point_cloud_range"], T, cap, overflow=overflow)
        o = oracle.OracleVoxelGenerator(c["voxel_size"], c["point_cloud_range"], T, cap, overflow)
        wv, wc, wn, woff = o.generate_batched(frames)
        gx, gy, gz = (int(x) for x in g.grid_size)
        out = None
        for rep in range(2):
            out = g.generate_bev(cat, offs, out=out)
            canvas, feats, coors, num, fvo = out
            torch.cuda.synchronize()
            assert np.array_equal(fvo.cpu().numpy(), woff)
            m = int(woff[-1])
            assert np.array_equal(coors[:m].cpu().numpy(), wc) and np.array_equal(num[:m].cpu().numpy(), wn)
            wf = oracle.mean_vfe(wv, wn)
            np.testing.assert_allclose(feats[:m].cpu().numpy(), wf, rtol=1e-6, atol=0)
            want = oracle.scatter_bev(wf, wc, len(sizes), gz, gy, gx)
            assert canvas.shape == (len(sizes), 4 * gz, gy, gx)
            np.testing.assert_allclose(canvas.cpu().numpy(), want, rtol=1e-6, atol=0)
        g.close()


@pytest.mark.parametrize("overflow", ["break", "continue"])
def test_cuda_graph_replay(pkg, oracle, synth, overflow):
    """The device forms enqueue only kernels (no host sync, no allocation of their own), so a call can
    be captured into a CUDA graph -- programmatic-dependent-launch edges included -- and replayed on
    new data in the same buffers: single frame (padded form) and a batch with the fused mean VFE."""
    g, o = _pair(pkg, oracle, KITTI, 5, 2500, overflow)
    frames = [synth.lidar_frame(20000, "kitti", seed=500 + i) for i in range(4)]
    buf = torch.empty((20000, 4), dtype=torch.float32, device="cuda")
    buf.copy_(torch.from_numpy(frames[0]))
    graph, (v, c, n, cnt) = g.capture(buf)
    for f in frames[1:] + frames[:1]:
        buf.copy_(torch.from_numpy(f))
        graph.replay()
        torch.cuda.synchronize()
        m = int(cnt.item())
        _eq((v[:m], c[:m], n[:m]), o.generate(f), "graph replay, single frame")
        # a plain call of another size on the same generator in between must not disturb the graph
        other = synth.lidar_frame(7000 + 1000 * m % 3000, "kitti", seed=77)
        _eq(g.generate(torch.from_numpy(other).cuda()), o.generate(other), "plain call between replays")
    # batched, out= buffers, fused mean
    cat = torch.empty((3 * 20000, 4), dtype=torch.float32, device="cuda")
    offs = np.asarray([0, 20000, 40000, 60000], np.int32)
    cat.copy_(torch.from_numpy(np.concatenate(frames[:3])))
    out = g.generate_batched(cat, offs, mean_channels=4, padded=True)
    torch.cuda.synchronize()
    graph2 = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph2):
        g.generate_batched(cat, offs, mean_channels=4, padded=True, out=out)
    for order in ([1, 2, 3], [3, 0, 2]):
        fr = [frames[i] for i in order]
        cat.copy_(torch.from_numpy(np.concatenate(fr)))
        graph2.replay()
        torch.cuda.synchronize()
        wv, wc, wn, woff = o.generate_batched(fr)
        assert np.array_equal(out[3].cpu().numpy(), woff)
        m = int(woff[-1])
        _eq((out[0][:m], out[1][:m], out[2][:m]), (wv, wc, wn), "graph replay, batch")
        np.testing.assert_allclose(out[4][:m].cpu().numpy(), oracle.mean_vfe(wv, wn), rtol=1e-6, atol=0)


def test_errors(pkg):
    g = pkg.VoxelGenerator([0.05, 0.05, 0.1], [0, -40, -3, 70.4, 40, 1], 5, 100)
    with pytest.raises(ValueError):
        g.generate(np.zeros((10, 2), np.float32))
    with pytest.raises(ValueError):
        g.generate(torch.zeros((10, 4)))            # CPU tensor
    with pytest.raises(ValueError):
        pkg.VoxelGenerator([0.05, 0.05, 0.1], [0, -40, -3, 70.4, 40, 1], 5, overflow="wrap")
    with pytest.raises(ValueError):
        pkg.VoxelGenerator([0.0, 0.05, 0.1], [0, -40, -3, 70.4, 40, 1], 5)
    with pytest.raises(RuntimeError):
        pkg.VoxelFeatureExtractorV3()(torch.zeros((3, 5, 4)), torch.ones(3, dtype=torch.int32))


@pytest.mark.parametrize("overflow", ["break", "continue"])
def test_frame_stats(pkg, oracle, synth, overflow):
    """bvx_frame_stats (what bench.py's `touched bytes` figure is made of): per frame the distinct cells among the points
    looked at, M, the break cut-off point and the processed prefix, against the same facts derived with numpy."""
    cap, T = 1200, 5
    g, o = _pair(pkg, oracle, KITTI, T, cap, overflow)
    frames = [synth.lidar_frame(n, "kitti", seed=700 + i) for i, n in enumerate([30000, 0, 9000, 2500])]
    got = g.generate_batched([torch.from_numpy(f).cuda() for f in frames])
    st = g.frame_stats(len(frames))
    want = o.generate_batched(frames)
    lo = np.asarray(KITTI["point_cloud_range"][:3], np.float32)
    hi = np.asarray(KITTI["point_cloud_range"][3:], np.float32)
    vs = np.asarray(KITTI["voxel_size"], np.float32)
    gs = np.round((hi - lo) / vs).astype(np.int64)
    assert np.array_equal(st["voxels"], np.diff(want[3]))
    for i, f in enumerate(frames):
        c = np.floor((f[:, :3] - lo) / vs).astype(np.int64)
        ok = np.all((c >= 0) & (c < gs), axis=1)
        key = (c[:, 2] * gs[1] + c[:, 1]) * gs[0] + c[:, 0]
        idx = np.nonzero(ok)[0]
        _, first = np.unique(key[idx], return_index=True)
        firsts = np.sort(idx[first])                      # point index of every cell's first point, in order
        n = f.shape[0]
        if overflow == "break" and len(firsts) > cap:
            assert st["cutoff"][i] == firsts[cap]         # the point that would open voxel number cap
            assert firsts[cap] < st["processed"][i] <= n and (st["processed"][i] % 512 == 0 or st["processed"][i] == n)
            assert st["cells"][i] == np.searchsorted(firsts, st["processed"][i])   # cells among the processed prefix
        else:
            assert st["processed"][i] == n and st["cutoff"][i] == n
            assert st["cells"][i] == len(firsts)
