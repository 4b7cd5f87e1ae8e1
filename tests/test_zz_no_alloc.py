"""bvx_scatter_bev_static -- the BEV scatter for callers that cannot allocate (no counterpart in the reference).
Called through the C ABI and compared with the existing entry point
bvx_scatter_bev on the same inputs: bit-equal (both run the same two kernels; only where the scratch map lives
differs), on two calls in a row with DIFFERENT inputs -- the second call finds the first call's entries in the
shared static map and must not be disturbed by them.  Collected last on purpose;
passes alone in a fresh process and after the other tests."""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


def _case(rng, b, c, d, ny, nx, m):
    cells = rng.choice(b * d * ny * nx, size=m, replace=False)
    bb, rem = np.divmod(cells, d * ny * nx)
    zz, rem = np.divmod(rem, ny * nx)
    yy, xx = np.divmod(rem, nx)
    coors = np.stack([bb, zz, yy, xx], axis=1).astype(np.int32)
    feats = rng.normal(0, 1, size=(m, c)).astype(np.float32)
    return torch.from_numpy(feats).cuda(), torch.from_numpy(coors).cuda()


def _call(lib, name, feats, coors, c, b, d, ny, nx):
    canvas = torch.full((b, c * d, ny, nx), float("nan"), dtype=torch.float32, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    rc = getattr(lib, name)(feats.data_ptr(), coors.data_ptr(), feats.shape[0], c, b, d, ny, nx, canvas.data_ptr(), st)
    return rc, canvas


def test_static_scatter_equals_scatter_bev_two_calls_in_a_row(pkg):
    lib = pkg._lib.load()
    rng = np.random.default_rng(77)
    # cfg3's canvas with 4 and 64 channels, a D > 1 canvas whose nx is not a multiple of 4, then a smaller call
    # whose cells were all used (with other voxel numbers) by the calls before it
    shapes = [(1, 4, 1, 496, 432, 12000), (1, 64, 1, 496, 432, 9000), (2, 3, 2, 37, 29, 700), (1, 4, 1, 496, 432, 300)]
    for (b, c, d, ny, nx, m) in shapes:
        feats, coors = _case(rng, b, c, d, ny, nx, m)
        rc0, want = _call(lib, "bvx_scatter_bev", feats, coors, c, b, d, ny, nx)
        rc1, got = _call(lib, "bvx_scatter_bev_static", feats, coors, c, b, d, ny, nx)
        torch.cuda.synchronize()
        assert rc0 == 0 and rc1 == 0, pkg._lib.load().bvx_last_error(None)
        assert torch.equal(got, want), (b, c, d, ny, nx, m)
        assert int((got != 0).sum().item()) > 0
    # m = 0: an all-zero canvas, no map involved
    empty_f = torch.empty((0, 4), dtype=torch.float32, device="cuda")
    empty_c = torch.empty((0, 4), dtype=torch.int32, device="cuda")
    rc, z = _call(lib, "bvx_scatter_bev_static", empty_f, empty_c, 4, 1, 1, 31, 29)
    torch.cuda.synchronize()
    assert rc == 0 and not bool(z.isnan().any()) and float(z.abs().sum().item()) == 0.0


def test_static_scatter_rejects_sizes_beyond_its_map(pkg):
    lib = pkg._lib.load()
    feats = torch.zeros((1, 1), dtype=torch.float32, device="cuda")
    coors = torch.zeros((1, 4), dtype=torch.int32, device="cuda")
    canvas = torch.zeros((16,), dtype=torch.float32, device="cuda")      # never written: the call must fail first
    st = torch.cuda.current_stream().cuda_stream
    rc = lib.bvx_scatter_bev_static(feats.data_ptr(), coors.data_ptr(), 1, 1, 1, 1, 1025, 1024, canvas.data_ptr(), st)
    assert rc == pkg._lib.BVX_ERR_WORKSPACE
    msg = lib.bvx_last_error(None).decode()
    assert "exceed" in msg and "1048576" in msg
    with pytest.raises(pkg._lib.WorkspaceError):
        pkg._lib.check(rc)
    # exactly at the limit is accepted (1024 x 1024 cells)
    big = torch.empty((1, 1, 1024, 1024), dtype=torch.float32, device="cuda")
    rc = lib.bvx_scatter_bev_static(feats.data_ptr(), coors.data_ptr(), 1, 1, 1, 1, 1024, 1024, big.data_ptr(), st)
    torch.cuda.synchronize()
    assert rc == 0 and float(big.abs().sum().item()) == 0.0
