"""Generate tests/golden/anchors_reference.npz by RUNNING THE REFERENCE's own
``sparse_sum_for_anchors_mask`` and ``fused_get_anchors_area``
(second/core/box_np_ops.py:776-781, :784-810) in this container.

Run from the repo root:  python tests/golden/make_golden_anchors.py
Needs /root/reference (read-only) and numba.  ``second.core.box_np_ops`` cannot be
imported without a GPU (it pulls in numba-CUDA kernels that compile at import time,
SURVEY.md section 0), so the two function definitions are cut out of the module's
syntax tree and compiled, decorators included, in a namespace that holds only
``np`` and ``numba`` -- the code that runs is the reference's, unmodified.  The
summed-area table in between is ``dense.cumsum(0).cumsum(1)`` as at
second/kittiviewer/viewer.py:397-398.
"""
import ast
import os

import numba
import numpy as np

REF_FILE = "/root/reference/second/core/box_np_ops.py"
HERE = os.path.dirname(os.path.abspath(__file__))
WANTED = ("sparse_sum_for_anchors_mask", "fused_get_anchors_area")


def reference_functions():
    src = open(REF_FILE).read()
    tree = ast.parse(src)
    nodes = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in WANTED]
    assert len(nodes) == len(WANTED)
    ns = {"np": np, "numba": numba}
    exec(compile(ast.Module(body=nodes, type_ignores=[]), REF_FILE, "exec"), ns)
    return ns["sparse_sum_for_anchors_mask"], ns["fused_get_anchors_area"]


def anchor_boxes(rng, n, rg, sizes):
    """(x0, y0, x1, y1) of boxes centred in (and a little beyond) the range."""
    cx = rng.uniform(rg[0] - 1.0, rg[3] + 1.0, n)
    cy = rng.uniform(rg[1] - 1.0, rg[4] + 1.0, n)
    wh = np.asarray(sizes, np.float64)[rng.integers(0, len(sizes), n)]
    bv = np.stack([cx - wh[:, 0] / 2, cy - wh[:, 1] / 2, cx + wh[:, 0] / 2, cy + wh[:, 1] / 2], 1)
    # the reference reads out of bounds for a box wholly outside the range: keep an overlap
    bv[:, 2] = np.maximum(bv[:, 2], rg[0] + 0.01)
    bv[:, 3] = np.maximum(bv[:, 3], rg[1] + 0.01)
    bv[:, 0] = np.minimum(bv[:, 0], rg[3] - 0.01)
    bv[:, 1] = np.minimum(bv[:, 1], rg[4] - 0.01)
    return bv.astype(np.float32)


def main():
    sparse_sum, fused = reference_functions()
    rng = np.random.default_rng(20261104)
    out = {}
    cases = {
        # tag: voxel_size, range, voxels, anchors
        "pillar": ([0.16, 0.16, 4.0], [0, -39.68, -3, 69.12, 39.68, 1], 3000, 1500),
        "second": ([0.2, 0.2, 0.4], [0, -40, -3, 70.4, 40, 1], 5000, 1500),
        "odd_nx": ([0.5, 0.25, 1.0], [-10, -10, -2, 25.5, 10, 2], 700, 400),
    }
    for tag, (vs, rg, m, a) in cases.items():
        vs = np.asarray(vs, np.float32)
        rg = np.asarray(rg, np.float32)
        grid = np.round((rg[3:] - rg[:3]) / vs).astype(np.int64)
        # distinct (z, y, x) voxels, several per pillar
        lin = rng.choice(int(grid.prod()), size=m, replace=False)
        z, r = np.divmod(lin, int(grid[1] * grid[0]))
        y, x = np.divmod(r, int(grid[0]))
        coors = np.stack([z, y, x], 1).astype(np.int32)
        bv = anchor_boxes(rng, a, rg, [(1.6, 3.9), (3.9, 1.6), (0.6, 0.8), (0.6, 1.76), (1.76, 0.6)])
        dense = sparse_sum(coors, tuple(grid[::-1][1:]))
        hist = dense.copy()
        dense = dense.cumsum(0)
        dense = dense.cumsum(1)
        area = fused(dense, bv, vs, rg, grid)
        out[f"{tag}_voxel_size"] = vs
        out[f"{tag}_range"] = rg
        out[f"{tag}_grid"] = grid
        out[f"{tag}_coors"] = coors
        out[f"{tag}_anchors_bv"] = bv
        out[f"{tag}_hist_nonzero"] = np.stack(np.nonzero(hist), 1).astype(np.int32)
        out[f"{tag}_hist_values"] = hist[np.nonzero(hist)].astype(np.float32)
        out[f"{tag}_area"] = area.astype(np.float32)
        print(tag, grid.tolist(), "voxels", m, "anchors", a, "area sum", float(area.sum()),
              "mask(>1)", int((area > 1).sum()))
    np.savez_compressed(os.path.join(HERE, "anchors_reference.npz"), **out)


if __name__ == "__main__":
    main()
