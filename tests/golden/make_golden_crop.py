"""Generate tests/golden/crop_reference.npz by RUNNING THE REFERENCE's own convex-polygon point
test -- ``points_in_convex_polygon_3d_jit`` / ``surface_equ_3d`` (second/core/geometry.py:84-141)
on surfaces built by ``corner_to_surfaces_3d_jit`` (second/core/box_np_ops.py:726-746), the chain
``remove_outside_points`` (box_np_ops.py:645-657) ends in.

Run from the repo root:  python tests/golden/make_golden_crop.py
Needs /root/reference (read-only) and numba.  geometry.py is loaded as a file (it only needs numpy
and numba); ``corner_to_surfaces_3d_jit`` is cut out of box_np_ops.py's syntax tree and compiled
with its decorator, because that module cannot be imported without a GPU.
The frustum corners are synthetic (a camera-like wedge in front of the sensor), ordered like the
reference's ``get_frustum`` output: near rectangle then far rectangle.
"""
import ast
import importlib.util
import os

import numba
import numpy as np

REF = "/root/reference/second/core"
HERE = os.path.dirname(os.path.abspath(__file__))


def reference_functions():
    spec = importlib.util.spec_from_file_location("ref_geometry", os.path.join(REF, "geometry.py"))
    geo = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(geo)
    src = open(os.path.join(REF, "box_np_ops.py")).read()
    nodes = [n for n in ast.parse(src).body
             if isinstance(n, ast.FunctionDef) and n.name == "corner_to_surfaces_3d_jit"]
    ns = {"np": np, "numba": numba}
    exec(compile(ast.Module(body=nodes, type_ignores=[]), "box_np_ops.py", "exec"), ns)
    return geo, ns["corner_to_surfaces_3d_jit"]


def wedge(near, far, half_w, half_h, yaw):
    """8 frustum corners ordered like the reference's get_frustum (box_np_ops.py:485-502): near
    rectangle then far rectangle, each (u0,v0), (u0,v1), (u1,v1), (u1,v0) in camera coordinates
    (x right, y down, z depth); then camera -> lidar axes (x = z_c, y = -x_c, z = -y_c) and a yaw."""
    cam = []
    for depth in (near, far):
        w, h = half_w * depth, half_h * depth
        cam += [[-w, -h, depth], [-w, h, depth], [w, h, depth], [w, -h, depth]]
    cam = np.asarray(cam, np.float64)
    lidar = np.stack([cam[:, 2], -cam[:, 0], -cam[:, 1]], 1)
    c, s = np.cos(yaw), np.sin(yaw)
    R = np.asarray([[c, -s, 0], [s, c, 0], [0, 0, 1]])
    return lidar @ R.T


def main():
    geo, corner_to_surfaces = reference_functions()
    rng = np.random.default_rng(20261104)
    out = {}
    for tag, (yaw, n) in {"front": (0.0, 6000), "left": (0.6, 6000)}.items():
        corners = wedge(0.5, 80.0, 0.9, 0.35, yaw)
        surfaces = corner_to_surfaces(corners[np.newaxis, ...])
        pts = np.concatenate([rng.uniform([-10, -45, -4], [75, 45, 3], (n, 3)),
                              rng.uniform(0, 1, (n, 1))], 1).astype(np.float32)
        # a few points exactly on a surface and at the apex side
        pts[:8, :3] = corners.astype(np.float32)
        mask = geo.points_in_convex_polygon_3d_jit(pts[:, :3], surfaces).reshape(-1)
        normal_vec, d = geo.surface_equ_3d(surfaces[:, :, :3, :])
        out[f"{tag}_points"] = pts
        out[f"{tag}_normal_vec"] = np.asarray(normal_vec[0], np.float64)
        out[f"{tag}_d"] = np.asarray(d[0], np.float64)
        out[f"{tag}_mask"] = mask
        print(tag, "inside", int(mask.sum()), "of", n)
        assert 0 < mask.sum() < n
    np.savez_compressed(os.path.join(HERE, "crop_reference.npz"), **out)


if __name__ == "__main__":
    main()
