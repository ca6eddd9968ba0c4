"""Deterministic synthetic inputs for the parity tests and the benchmark.

Producers of INPUTS only (volumes, cameras, transfer functions); nothing here is on the hot
path.  Each generator cites the reference code whose semantics it follows.
"""
from __future__ import annotations

import math
from typing import Tuple

import numpy as np


# ---- volumes -------------------------------------------------------------------------------
def marschner_lobb(n: int, fM: float = 6.0, alpha: float = 0.25) -> np.ndarray:
    """Marschner-Lobb test signal sampled on [-1,1]^3 inclusive, shape (1,n,n,n) fp32, index
    order [b][x][y][z] (renderer/volume_implicit.cpp:28-42 evaluated in double,
    renderer/volume.cpp:369-386 for the sampling grid)."""
    scale = np.float32(2.0) / np.float32(n - 1)
    c = (np.float32(-1.0) + np.arange(n, dtype=np.float32) * scale).astype(np.float64)
    x, y, z = np.meshgrid(c, c, c, indexing="ij")
    r = np.sqrt(x * x + y * y)
    pr = np.cos(2 * np.pi * fM * np.cos(np.pi * r / 2))
    num = (1 - np.sin(np.pi * z / 2)) + alpha * (1 + pr)
    return (num / (2 * (1 + alpha))).astype(np.float32)[None]


def _value_noise(n: int, rng: np.random.RandomState, octaves=(8, 16, 32, 64)):
    """4-octave value noise in [0,1]: random lattices upsampled trilinearly (torch, CPU)."""
    import torch
    import torch.nn.functional as F

    out = torch.zeros(1, 1, n, n, n, dtype=torch.float32)
    amp, total = 1.0, 0.0
    for res in octaves:
        lattice = torch.from_numpy(rng.rand(1, 1, res, res, res).astype(np.float32))
        out += amp * F.interpolate(lattice, size=(n, n, n), mode="trilinear", align_corners=True)
        total += amp
        amp *= 0.5
    return (out / total)[0, 0]


def skull(n: int, seed: int = 7):
    """Builder-defined skull-like density on [-1,1]^3: a thin dense ellipsoidal
    shell, noisy soft tissue inside, a dense 'jaw' torus; ~40 % of the box is exactly empty.
    Returns a torch CPU tensor (1,n,n,n) fp32, index order [b][x][y][z]."""
    import torch

    rng = np.random.RandomState(seed)
    c = torch.linspace(-1.0, 1.0, n, dtype=torch.float32)
    x, y, z = torch.meshgrid(c, c, c, indexing="ij")
    r_e = torch.sqrt((x / 1.0) ** 2 + (y / 0.8) ** 2 + (z / 0.9) ** 2)

    def shell(r, r0, w):
        return torch.exp(-((r - r0) / w) ** 2)

    def smoothstep(e0, e1, v):
        t = ((v - e0) / (e1 - e0)).clamp(0, 1)
        return t * t * (3 - 2 * t)

    fbm = _value_noise(n, rng)
    # jaw: torus of major radius 0.45 in the plane y = -0.45, minor radius given by the shell
    q = torch.sqrt(x * x + z * z) - 0.45
    r_j = torch.sqrt(q * q + (y + 0.45) ** 2)
    d = (0.9 * shell(r_e, 0.62, 0.05)
         + 0.35 * smoothstep(0.0, 0.1, 0.58 - r_e) * (0.5 + 0.5 * fbm)
         + 0.6 * shell(r_j, 0.0, 0.08))
    d = d.clamp(0, 1)
    d = torch.where(d < 1e-3, torch.zeros_like(d), d)  # exact zeros outside the head
    return d[None].contiguous()


# ---- cameras ---------------------------------------------------------------------------------
def fibonacci_sphere(N: int) -> Tuple[np.ndarray, np.ndarray]:
    """(pitch/latitude, yaw/longitude) of N points on a sphere, pytests/diffdvr/utils.py:76-89."""
    gr = (np.sqrt(5.0) + 1.0) / 2.0
    ga = (2 - gr) * (2 * np.pi)
    i = np.arange(1, N + 1, dtype=np.float32)
    lat = np.arcsin(-1 + 2 * i / (N + 1))
    lon = np.remainder(ga * i, 2 * np.pi)
    return lat.astype(np.float32), lon.astype(np.float32)


def _normalize(v):
    return v / np.linalg.norm(v, axis=-1, keepdims=True)


def viewport_from_look_at(origin, look_at, up) -> np.ndarray:
    """(B,3,3) = [eye, right, up] rows, renderer/camera.cpp:131-150."""
    origin, look_at, up = (np.asarray(a, dtype=np.float32).reshape(-1, 3) for a in (origin, look_at, up))
    front = _normalize(look_at - origin)
    right = _normalize(np.cross(front, up))
    up2 = _normalize(np.cross(right, front))
    return np.stack([np.broadcast_to(origin, front.shape), right, up2], axis=1).astype(np.float32)


def viewport_from_sphere(center, yaw, pitch, distance, orientation: str = "Ym") -> np.ndarray:
    """renderer/camera.cpp:152-207 (Orientation tables :68-87)."""
    names = ["Xp", "Xm", "Yp", "Ym", "Zp", "Zm"]
    ups = [(1, 0, 0), (-1, 0, 0), (0, 1, 0), (0, -1, 0), (0, 0, 1), (0, 0, -1)]
    perms = [(2, -1, -3), (-2, 1, 3), (1, 2, 3), (-1, -2, -3), (-3, -1, 2), (3, 1, -2)]
    inv_yaw = [True, False, False, True, False, True]
    o = names.index(orientation)
    yaw = np.asarray(yaw, dtype=np.float32).reshape(-1) * (1.0 if inv_yaw[o] else -1.0)
    pitch = np.asarray(pitch, dtype=np.float32).reshape(-1) * -1.0  # OrientationInvertPitch all false
    distance = np.broadcast_to(np.asarray(distance, dtype=np.float32).reshape(-1), yaw.shape)
    pos = [np.cos(pitch) * np.cos(yaw) * distance, np.sin(pitch) * distance,
           np.cos(pitch) * np.sin(yaw) * distance]
    pos2 = [pos[abs(p) - 1] * (1 if p > 0 else -1) for p in perms[o]]
    origin = np.stack(pos2, axis=1)
    center = np.asarray(center, dtype=np.float32).reshape(-1, 3)
    upv = np.broadcast_to(np.asarray(ups[o], dtype=np.float32), origin.shape)
    return viewport_from_look_at(origin + center, np.broadcast_to(center, origin.shape), upv)


def generate_rays(viewport: np.ndarray, fov_y: float, W: int, H: int):
    """Per-pixel rays of a reference-frame camera, renderer/camera_impl.cu:17-37,68-77.
    Returns (ray_start, ray_dir) each (B,H,W,3) fp32."""
    vp = np.asarray(viewport, dtype=np.float32)
    fov_x = np.float32(fov_y) * np.float32(W) / np.float32(H)
    tan_x, tan_y = np.float32(math.tan(fov_x / 2)), np.float32(math.tan(fov_y / 2))
    eye, right, up = vp[:, 0], vp[:, 1], vp[:, 2]
    front = np.cross(up, right)
    fx = (2 * (np.arange(W, dtype=np.float32) + 0.5) / np.float32(W) - 1).astype(np.float32)
    fy = (2 * (np.arange(H, dtype=np.float32) + 0.5) / np.float32(H) - 1).astype(np.float32)
    d = (front[:, None, None, :] + (fx * tan_x)[None, None, :, None] * right[:, None, None, :]
         + (fy * tan_y)[None, :, None, None] * up[:, None, None, :])
    d = d / np.sqrt((d * d).sum(-1, keepdims=True))
    start = np.broadcast_to(eye[:, None, None, :], d.shape)
    return np.ascontiguousarray(start, dtype=np.float32), np.ascontiguousarray(d, dtype=np.float32)


def inverse_view_projection(origin, look_at, up, fov_degrees, W, H, near=0.1, far=10.0) -> np.ndarray:
    """Row-major 4x4 inverse view-projection matrix, renderer/camera.cpp:45-127
    (glm::lookAtLH, perspectiveFovLH_ZO, inverse; the final glm::transpose makes the memory
    rows equal to the math rows, which is what the kernel's matmul expects)."""
    e, c, u = (np.asarray(a, dtype=np.float64) for a in (origin, look_at, up))
    u = u / np.linalg.norm(u)
    f = (c - e) / np.linalg.norm(c - e)
    s = np.cross(u, f)
    s /= np.linalg.norm(s)
    uu = np.cross(f, s)
    V = np.array([[s[0], s[1], s[2], -s.dot(e)], [uu[0], uu[1], uu[2], -uu.dot(e)],
                  [f[0], f[1], f[2], -f.dot(e)], [0, 0, 0, 1]], dtype=np.float64)
    fov = math.radians(fov_degrees)
    h = math.cos(0.5 * fov) / math.sin(0.5 * fov)
    w = h * H / W
    P = np.array([[w, 0, 0, 0], [0, h, 0, 0], [0, 0, far / (far - near), -(far * near) / (far - near)],
                  [0, 0, 1, 0]], dtype=np.float64)
    return np.linalg.inv(P @ V).astype(np.float32)


# ---- transfer functions ------------------------------------------------------------------------
def tf_linear_two_peaks(opacity_scale: float = 20.0) -> np.ndarray:
    """Piecewise-linear TF (1,6,5): r,g,b,absorption,position with two opacity peaks."""
    pts = [
        (0.0, 0.0, 0.0, 0.0, 0.00),
        (0.9, 0.3, 0.1, 0.0, 0.20),
        (0.9, 0.5, 0.2, 0.6, 0.35),
        (0.2, 0.4, 0.9, 0.05, 0.55),
        (1.0, 1.0, 0.9, 1.0, 0.80),
        (1.0, 1.0, 1.0, 1.0, 1.00),
    ]
    tf = np.array(pts, dtype=np.float32)
    tf[:, 3] *= opacity_scale
    return tf[None]


def tf_texture_from_linear(tf_linear: np.ndarray, R: int = 64) -> np.ndarray:
    """Texture TF (1,R,4) sampled at texel centres from a piecewise-linear TF."""
    pts = tf_linear[0]
    d = (np.arange(R, dtype=np.float32) + 0.5) / R
    out = np.stack([np.interp(d, pts[:, 4], pts[:, c]) for c in range(4)], axis=1)
    return out.astype(np.float32)[None]


def tf_gaussian(R: int = 4) -> np.ndarray:
    """Sum-of-Gaussians TF (1,R,6): r,g,b,absorption,mean,sigma (cf. pytests/diffdvr/settings.py:147-156)."""
    mus = np.linspace(0.25, 0.85, R)
    cols = [(0.9, 0.2, 0.1), (0.2, 0.8, 0.3), (0.2, 0.3, 0.9), (0.95, 0.95, 0.8)]
    ops = np.linspace(30.0, 100.0, R)
    rows = [(*cols[i % 4], ops[i], mus[i], 0.04) for i in range(R)]
    return np.array(rows, dtype=np.float32)[None]


def tf_identity(scale_absorption: float = 1.0, scale_color: float = 1.0) -> np.ndarray:
    return np.array([[[scale_absorption, scale_color]]], dtype=np.float32)


def default_box(n: int):
    """boxSize=(1,1,1), boxMin=-0.5-0.5/(n-1) (pytests/diffdvr/settings.py:170-173)."""
    m = -0.5 - 0.5 / (n - 1)
    return (m, m, m), (1.0, 1.0, 1.0)
