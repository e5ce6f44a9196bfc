"""Seeded synthetic scenes and camera trajectories for the BASELINE.json configs.

Host-side helpers only; nothing here is on the render path.
"""
from __future__ import annotations

import numpy as np

from .util import Camera, look_at
from .util_gau import GaussianData


def synthetic_gaussians(n: int, seed: int = 0, sh_dim: int = 48) -> GaussianData:
    """xyz 70 % N(0, 2^2) + 30 % U[-8, 8]^3; log-scale N(ln 0.02, 0.6^2) clipped to
    [ln 1e-3, ln 0.5]; rotation = normalised N(0,1)^4; opacity = sigmoid(N(0, 2^2));
    SH DC N(0, 1), higher bands N(0, 0.15^2).  All float32."""
    rng = np.random.default_rng(seed)
    n_gauss = int(round(0.7 * n))
    xyz = np.empty((n, 3), dtype=np.float32)
    xyz[:n_gauss] = rng.normal(0.0, 2.0, size=(n_gauss, 3)).astype(np.float32)
    xyz[n_gauss:] = rng.uniform(-8.0, 8.0, size=(n - n_gauss, 3)).astype(np.float32)
    perm = rng.permutation(n)
    xyz = xyz[perm]
    log_s = rng.normal(np.log(0.02), 0.6, size=(n, 3))
    log_s = np.clip(log_s, np.log(1e-3), np.log(0.5))
    scale = np.exp(log_s).astype(np.float32)
    rot = rng.normal(size=(n, 4))
    rot = (rot / np.linalg.norm(rot, axis=-1, keepdims=True)).astype(np.float32)
    opacity = (1.0 / (1.0 + np.exp(-rng.normal(0.0, 2.0, size=(n, 1))))).astype(np.float32)
    sh = np.empty((n, sh_dim), dtype=np.float32)
    sh[:, :3] = rng.normal(0.0, 1.0, size=(n, 3)).astype(np.float32)
    if sh_dim > 3:
        sh[:, 3:] = rng.normal(0.0, 0.15, size=(n, sh_dim - 3)).astype(np.float32)
    return GaussianData(xyz, rot, scale, opacity, sh)


def default_view(w: int, h: int, eye=(0.0, 0.0, 10.0), target=(0.0, 0.0, 0.0), up=(0.0, -1.0, 0.0)):
    """Single benchmark view: reference default fovy, eye (0,0,10) -> origin,
    up (0,-1,0).  Returns (camera, pose dict) ready for update_camera_pose(camera, True, pose)."""
    cam = Camera(h, w)
    eye = np.asarray(eye, dtype=np.float32)
    target = np.asarray(target, dtype=np.float32)
    front = target - eye
    front = front / np.linalg.norm(front)
    pose = {
        "camera_front": front.astype(np.float32),
        "camera_up": np.asarray(up, dtype=np.float32),
        "camera_position": eye,
        "camera_view": look_at(eye, target, up),
    }
    return cam, pose


def _catmull_rom_closed(points: np.ndarray, samples: int) -> np.ndarray:
    k = len(points)
    t = np.linspace(0.0, k, samples, endpoint=False)
    i1 = np.floor(t).astype(int) % k
    u = (t - np.floor(t))[:, None]
    p0, p1, p2, p3 = points[(i1 - 1) % k], points[i1], points[(i1 + 1) % k], points[(i1 + 2) % k]
    return 0.5 * ((2 * p1) + (-p0 + p2) * u + (2 * p0 - 5 * p1 + 4 * p2 - p3) * u ** 2 +
                  (-p0 + 3 * p1 - 3 * p2 + p3) * u ** 3)


def trajectory_poses(rows: np.ndarray, n_poses: int, scale: float = 3.0):
    """Config 5 trajectory: closed Catmull-Rom loop through the rows of the viewer's
    camera_data.csv (front xyz, up xyz, position xyz — writer /root/reference/main.py:422-434),
    positions scaled about the origin, resampled to n_poses.  Returns
    (views (n,4,4) f32 maths layout, campos (n,3) f32)."""
    rows = np.asarray(rows, dtype=np.float64).reshape(-1, 9)
    front = _catmull_rom_closed(rows[:, 0:3], n_poses)
    pos = _catmull_rom_closed(rows[:, 6:9] * scale, n_poses)
    up = np.array([0.0, -1.0, 0.0])
    views = np.empty((n_poses, 4, 4), dtype=np.float32)
    for i in range(n_poses):
        f = front[i] / max(np.linalg.norm(front[i]), 1e-9)
        views[i] = look_at(pos[i], pos[i] + f, up)
    return views, pos.astype(np.float32)


def orbit_poses(n_poses: int, radius: float = 10.0, height: float = 2.0):
    """Fallback trajectory that needs no fixture: a circular orbit looking at the origin."""
    ang = np.linspace(0.0, 2 * np.pi, n_poses, endpoint=False)
    pos = np.stack([radius * np.sin(ang), np.full_like(ang, height), radius * np.cos(ang)], axis=1)
    views = np.stack([look_at(p, (0, 0, 0), (0, -1, 0)) for p in pos]).astype(np.float32)
    return views, pos.astype(np.float32)
