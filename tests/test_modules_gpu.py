"""GPU tests of the module mirrors beyond the mean VFE: the learned VFEs and the
camera -> BEV index scatter (row 4) against fixtures produced by the reference's own classes / expressions
(tests/golden/make_golden_vfe.py), plus the autograd behaviour of the parameter-free modules."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")

CASES = {
    "v1_norm": ("VoxelFeatureExtractor", dict(num_input_features=4, use_norm=True, num_filters=[32, 128], with_distance=False)),
    "v1_dist": ("VoxelFeatureExtractor", dict(num_input_features=4, use_norm=False, num_filters=[16, 64], with_distance=True)),
    "v2_three": ("VoxelFeatureExtractorV2", dict(num_input_features=5, use_norm=True, num_filters=[16, 32, 64], with_distance=True)),
}


@pytest.mark.parametrize("tag", sorted(CASES))
def test_learned_vfe_against_reference_classes(pkg, golden_dir, tag):
    """Same parameters (the reference's state dict loads with strict=True), same outputs: the augmentation and
    the mask before the first layer to 1e-6, the voxel-wise features to 1e-4 (CPU vs GPU GEMM rounding)."""
    g = np.load(os.path.join(golden_dir, "vfe_learned_reference.npz"))
    cls, kw = CASES[tag]
    net = getattr(pkg, cls)(**kw)
    sd = {k[len(tag) + 4:]: torch.from_numpy(g[k]) for k in g.files if k.startswith(tag + "_sd_")}
    net.load_state_dict(sd, strict=True)
    net = net.cuda().eval()
    vox, num = torch.from_numpy(g[tag + "_voxels"]).cuda(), torch.from_numpy(g[tag + "_num"]).cuda()
    aug, mask = pkg.vfe_augment(vox, num, kw["with_distance"])
    np.testing.assert_allclose(aug.cpu().numpy(), g[tag + "_aug"], rtol=1e-6, atol=1e-6)
    assert np.array_equal(mask.cpu().numpy() != 0, g[tag + "_mask"])
    assert np.array_equal(pkg.get_paddings_indicator(num, vox.shape[1]).cpu().numpy(), g[tag + "_mask"])
    with torch.no_grad():
        y = net(vox, num)
    np.testing.assert_allclose(y.cpu().numpy(), g[tag + "_out"], rtol=1e-4, atol=1e-4)


def test_masked_max_and_augment_match_torch_and_backpropagate(pkg):
    rng = np.random.default_rng(5)
    m, t, f, c = 513, 7, 4, 33
    num = torch.from_numpy(rng.integers(1, t + 1, size=m).astype(np.int32)).cuda()
    x = torch.from_numpy(rng.normal(0, 1, size=(m, t, c)).astype(np.float32)).cuda().requires_grad_(True)
    mask = (torch.arange(t, device="cuda").view(1, -1) < num.view(-1, 1)).float().unsqueeze(-1)
    want = (x * mask).max(dim=1)[0]
    got = pkg.masked_max(x, num)
    assert torch.equal(got, want)                              # multiply-then-max, bit for bit
    gw = torch.autograd.grad(want.sum(), x)[0]
    gg = torch.autograd.grad(got.sum(), x)[0]
    assert torch.allclose(gg, gw)
    vox = torch.from_numpy(rng.normal(0, 5, size=(m, t, f)).astype(np.float32)).cuda()
    vox = (vox * mask).requires_grad_(True)
    for dist in (False, True):
        mean = vox[:, :, :3].sum(dim=1, keepdim=True) / num.float().view(-1, 1, 1)
        parts = [vox, vox[:, :, :3] - mean] + ([torch.norm(vox[:, :, :3], 2, 2, keepdim=True)] if dist else [])
        want = torch.cat(parts, dim=-1)
        got, _ = pkg.vfe_augment(vox, num, dist)
        assert torch.allclose(got, want, rtol=1e-6, atol=1e-6)
        w = torch.from_numpy(rng.normal(0, 1, size=tuple(want.shape)).astype(np.float32)).cuda()
        gw = torch.autograd.grad((want * w * mask).sum(), vox)[0]
        gg = torch.autograd.grad((got * w * mask).sum(), vox)[0]
        assert torch.allclose(gg, gw, rtol=1e-4, atol=1e-5)


@pytest.mark.parametrize("tag", ["fusion", "dup"])
def test_index_scatter_against_reference_expression(pkg, golden_dir, tag):
    """rpn.py:957-962 for the whole batch in one call; the fixture's targets collide (several pixels per BEV cell):
    the last pixel wins, as in the reference's sequential assignment."""
    g = np.load(os.path.join(golden_dir, "index_scatter_reference.npz"))
    src = torch.from_numpy(g[tag + "_src"]).cuda()
    idxs = torch.from_numpy(g[tag + "_idxs"]).cuda()
    want = g[tag + "_out"]
    ny, nx = want.shape[2:]
    pts = g[tag + "_points"]
    got = pkg.scatter_index_bev(src, idxs, ny, nx, src_idx=pts)
    assert np.array_equal(got.cpu().numpy(), want)
    # twice on the same stream (the map is reused), and through the implicit mesh when the points are one
    assert np.array_equal(pkg.scatter_index_bev(src, idxs, ny, nx, src_idx=pts).cpu().numpy(), want)
    s0, s1 = src.shape[2:]
    order = np.arange(s0 * s1)
    if np.array_equal(pts, np.stack([order // s1, order % s1], axis=1)):
        assert np.array_equal(pkg.scatter_index_bev(src, idxs, ny, nx).cpu().numpy(), want)


def test_parameter_free_modules_keep_dtype_and_gradients(pkg):
    rng = np.random.default_rng(9)
    m, t, f = 300, 5, 4
    num = torch.from_numpy(rng.integers(1, t + 1, size=m).astype(np.int32)).cuda()
    mask = (torch.arange(t, device="cuda").view(1, -1) < num.view(-1, 1)).float().unsqueeze(-1)
    vox = (torch.from_numpy(rng.normal(0, 5, size=(m, t, f)).astype(np.float32)).cuda() * mask).requires_grad_(True)
    # mean VFE: gradient = grad / num broadcast over T
    w = torch.from_numpy(rng.normal(0, 1, size=(m, 4)).astype(np.float32)).cuda()
    got = pkg.VoxelFeatureExtractorV3()(vox, num)
    want = vox[:, :, :4].sum(dim=1) / num.float().view(-1, 1)
    assert torch.allclose(got, want, rtol=1e-6, atol=1e-6)
    assert torch.allclose(torch.autograd.grad((got * w).sum(), vox)[0], torch.autograd.grad((want * w).sum(), vox)[0])
    # scatter: dtype kept, rows outside the canvas skipped in forward AND backward
    b, d, ny, nx, c = 2, 1, 12, 10, 6
    cells = rng.choice(b * ny * nx, size=50, replace=False)
    co = np.stack([cells // (ny * nx), np.zeros(50, np.int64), (cells // nx) % ny, cells % nx], axis=1).astype(np.int32)
    co[7] = [0, 0, ny, 3]                   # outside
    co[9] = [-1, 0, 2, 2]                   # outside
    feats = torch.from_numpy(rng.normal(0, 1, size=(50, c)).astype(np.float32)).cuda().requires_grad_(True)
    canvas = pkg.scatter_bev(feats, torch.from_numpy(co).cuda(), b, d, ny, nx)
    ok = np.ones(50, bool); ok[[7, 9]] = False
    want = np.zeros((b, c, ny, nx), np.float32)
    for i in np.nonzero(ok)[0]:
        want[co[i, 0], :, co[i, 2], co[i, 3]] = feats.detach().cpu().numpy()[i]
    assert np.array_equal(canvas.detach().cpu().numpy(), want)
    gw = torch.from_numpy(rng.normal(0, 1, size=(b, c, ny, nx)).astype(np.float32)).cuda()
    gf = torch.autograd.grad((canvas * gw).sum(), feats)[0].cpu().numpy()
    for i in range(50):
        exp = gw[co[i, 0], :, co[i, 2], co[i, 3]].cpu().numpy() if ok[i] else np.zeros(c, np.float32)
        assert np.allclose(gf[i], exp)
    half = pkg.scatter_bev(feats.detach().half(), torch.from_numpy(co).cuda(), b, d, ny, nx)
    assert half.dtype == torch.float16
