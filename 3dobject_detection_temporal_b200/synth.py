"""Deterministic synthetic point clouds and the frozen measurement configs.

Spinning-lidar model of SURVEY.md section 8(d) / BASELINE.md section 2: R rings with
elevations linearly spaced in [e0, e1], N/R azimuths per ring ~ U[-pi, pi); range
= min(ground hit at sensor height h, Gamma(2, 12 m), 80 m) * (1 + N(0, 0.002));
intensity ~ U[0, 1); rows shuffled (``shuffle_points: true`` in the reference's
train configs, second/configs/onestage/server_lidaronly.fhd.onestage.config:109)
or left in ring order (eval, :183).  fp32, no NaN/Inf.

The on-disk point format these mimic is ``float32.reshape(-1, 4)``
(second/create_data.py:25-26).
"""
import math

import numpy as np

SENSORS = {
    # name: (rings, elevation range in degrees, sensor height in m)
    "kitti": (64, (-24.8, 2.0), 1.73),
    "nuscenes": (32, (-30.67, 10.67), 1.84),
    "waymo": (64, (-17.6, 2.4), 2.0),
}


def lidar_frame(n, sensor="kitti", seed=0, shuffle=True):
    """One [n, 4] float32 frame (x, y, z, intensity) in the sensor frame."""
    rings, (e0, e1), h = SENSORS[sensor]
    rng = np.random.default_rng(seed)
    ring = (np.arange(n, dtype=np.int64) * rings) // max(n, 1)
    elev = np.deg2rad(np.linspace(e0, e1, rings))[ring]
    az = rng.uniform(-math.pi, math.pi, size=n)
    with np.errstate(divide="ignore", invalid="ignore"):
        ground = np.where(elev < 0, h / np.sin(-elev), np.inf)
    obj = rng.gamma(2.0, 12.0, size=n)
    r = np.minimum(np.minimum(ground, obj), 80.0) * (1.0 + rng.normal(0.0, 0.002, size=n))
    pts = np.empty((n, 4), dtype=np.float32)
    pts[:, 0] = r * np.cos(elev) * np.cos(az)
    pts[:, 1] = r * np.cos(elev) * np.sin(az)
    pts[:, 2] = r * np.sin(elev)
    pts[:, 3] = rng.uniform(0.0, 1.0, size=n)
    if shuffle:
        rng.shuffle(pts, axis=0)
    return np.ascontiguousarray(pts)


def sweep_transform(s, yaw_deg_per_sweep=0.5, tx_per_sweep=0.5):
    """Sweep->key-frame rigid transform used by cfg2/cfg5: yaw 0.5 deg * s,
    translation (0.5 * s, 0, 0) m.  s = 0 is the key frame (identity)."""
    a = math.radians(yaw_deg_per_sweep * s)
    R = np.array([[math.cos(a), -math.sin(a), 0.0],
                  [math.sin(a), math.cos(a), 0.0],
                  [0.0, 0.0, 1.0]], dtype=np.float32)
    t = np.array([tx_per_sweep * s, 0.0, 0.0], dtype=np.float32)
    return R, t


def lidar_sweeps(k, n_per_sweep, sensor, seed0, dt_per_sweep, shuffle=True):
    """List of (points, R, t, dt); the key frame has R = t = None."""
    out = []
    for s in range(k):
        pts = lidar_frame(n_per_sweep, sensor, seed0 + s, shuffle)
        if s == 0:
            out.append((pts, None, None, 0.0))
        else:
            R, t = sweep_transform(s)
            out.append((pts, R, t, float(np.float32(dt_per_sweep * s))))
    return out


# BASELINE.md section 2 -- frozen measurement configs
CONFIGS = {
    "cfg1": dict(sensor="kitti", n=120000, sweeps=1, frames=1, f_in=4,
                 point_cloud_range=[0, -40, -3, 70.4, 40, 1], voxel_size=[0.05, 0.05, 0.1],
                 max_num_points=5, max_voxels=20000, seed0=0, dt=0.0),
    "cfg2": dict(sensor="nuscenes", n=34000, sweeps=5, frames=1, f_in=4,
                 point_cloud_range=[-51.2, -51.2, -5, 51.2, 51.2, 3], voxel_size=[0.1, 0.1, 0.2],
                 max_num_points=10, max_voxels=60000, seed0=100, dt=0.05),
    "cfg3": dict(sensor="kitti", n=120000, sweeps=1, frames=1, f_in=4,
                 point_cloud_range=[0, -39.68, -3, 69.12, 39.68, 1], voxel_size=[0.16, 0.16, 4],
                 max_num_points=100, max_voxels=12000, seed0=0, dt=0.0),
    "cfg4": dict(sensor="kitti", n=120000, sweeps=1, frames=64, f_in=4,
                 point_cloud_range=[0, -40, -3, 70.4, 40, 1], voxel_size=[0.05, 0.05, 0.1],
                 max_num_points=5, max_voxels=20000, seed0=0, dt=0.0),
    "cfg5": dict(sensor="waymo", n=180000, sweeps=3, frames=1, f_in=4,
                 point_cloud_range=[-75.2, -75.2, -2, 75.2, 75.2, 4], voxel_size=[0.1, 0.1, 0.15],
                 max_num_points=5, max_voxels=150000, seed0=200, dt=0.1),
}


def make_frames(cfg_name, frames=None, shuffle=True, seed_offset=0):
    """Frames of a single-sweep config: list of [n, 4] arrays, seeds seed0 + i."""
    c = CONFIGS[cfg_name]
    assert c["sweeps"] == 1
    nb = c["frames"] if frames is None else frames
    return [lidar_frame(c["n"], c["sensor"], c["seed0"] + seed_offset + i, shuffle)
            for i in range(nb)]


def make_sweeps(cfg_name, shuffle=True):
    c = CONFIGS[cfg_name]
    return lidar_sweeps(c["sweeps"], c["n"], c["sensor"], c["seed0"], c["dt"], shuffle)
