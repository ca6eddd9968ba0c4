"""Seeded inputs for Camera::generateRays (forward + backward) and Renderer::compareToImage and the
one routine that runs ANY library exporting those C-ABI entry points on them: the reference-built
shim (oracle/_ref) when generating the fixtures tests/golden/ref_camera_*.npz, the CPU port and the
CUDA product when checking."""
from __future__ import annotations

import math

import numpy as np

SEED = 77


def viewport_from_lookat(origin, look_at, up):
    """numpy restatement of Camera::viewportFromLookAt (renderer/camera.cpp:118-139)."""
    def norm(v):
        return v / np.maximum(np.linalg.norm(v, axis=1, keepdims=True), 1e-12)
    front = norm(look_at - origin)
    right = norm(np.cross(front, up, axis=1))
    up2 = norm(np.cross(right, front, axis=1))
    return np.stack([origin, right, up2], axis=1)


def make_viewports(B, rng, dtype=np.float32):
    origin = rng.uniform(-1, 1, (B, 3))
    origin = 1.6 * origin / np.linalg.norm(origin, axis=1, keepdims=True)
    look_at = rng.uniform(-0.1, 0.1, (B, 3))
    up = np.tile(np.array([[0.0, 0.0, 1.0]]), (B, 1)) + rng.uniform(-0.2, 0.2, (B, 3))
    return np.ascontiguousarray(viewport_from_lookat(origin, look_at, up).astype(dtype))


def ray_cases():
    """name -> dict(B, W, H, fovY); W % 4 == 0 exercises the packed 128-bit path of the CUDA kernels."""
    return {
        "rays_packed": dict(B=3, W=24, H=10, fovY=math.radians(45.0)),
        "rays_odd": dict(B=2, W=13, H=9, fovY=math.radians(30.0)),
        "rays_single": dict(B=1, W=1, H=1, fovY=math.radians(60.0)),
    }


def compare_cases():
    return {
        "compare_d1": dict(B=2, W=12, H=7, D=1),
        "compare_d6": dict(B=1, W=9, H=5, D=6),
    }


def run_rays(lib, case, dtype=np.float32, wrap=None, unwrap=None, sync=None):
    """wrap: numpy -> tensor the library can take (identity for host libraries, .cuda() for the product)."""
    wrap = wrap or (lambda a: a)
    unwrap = unwrap or (lambda t: t)
    rng = np.random.RandomState(SEED)
    B, W, H, fovY = case["B"], case["W"], case["H"], case["fovY"]
    vp = make_viewports(B, rng, dtype)
    adj_start = rng.standard_normal((B, H, W, 3)).astype(dtype)
    adj_dir = rng.standard_normal((B, H, W, 3)).astype(dtype)
    t_vp = wrap(vp)
    start, direction = wrap(np.zeros((B, H, W, 3), dtype)), wrap(np.zeros((B, H, W, 3), dtype))
    lib.generate_rays(t_vp, fovY, start, direction)
    adj_vp = wrap(np.full((B, 3, 3), 123.0, dtype))  # must be overwritten, not accumulated
    lib.generate_rays_adjoint(t_vp, fovY, wrap(adj_start), wrap(adj_dir), adj_vp)
    if sync:
        sync()
    return dict(viewport=vp, ray_start=unwrap(start), ray_dir=unwrap(direction), adj_viewport=unwrap(adj_vp),
                adj_ray_start=adj_start, adj_ray_dir=adj_dir)


def run_compare(lib, case, dtype=np.float32, wrap=None, unwrap=None, sync=None):
    wrap = wrap or (lambda a: a)
    unwrap = unwrap or (lambda t: t)
    rng = np.random.RandomState(SEED + 1)
    B, W, H, D = case["B"], case["W"], case["H"], case["D"]
    color = rng.uniform(0, 1, (B, H, W, 4)).astype(dtype)
    reference = rng.uniform(0, 1, (B, H, W, 4)).astype(dtype)
    grads = rng.standard_normal((B, H, W, D, 4)).astype(dtype)
    tc, tg, tr = wrap(color), wrap(grads), wrap(reference)
    diff = wrap(np.full((B, H, W, 1), 9.0, dtype))
    gout = wrap(np.full((B, H, W, D), 9.0, dtype))
    lib.compare_to_image(tc, tg, tr, diff, gout, False)
    diff_r = wrap(np.full((1,), 9.0, dtype))
    gout_r = wrap(np.full((D,), 9.0, dtype))
    lib.compare_to_image(tc, tg, tr, diff_r, gout_r, True)
    if sync:
        sync()
    return dict(difference=unwrap(diff), gradients=unwrap(gout), difference_reduced=unwrap(diff_r),
                gradients_reduced=unwrap(gout_r))


def preshade_cases():
    """name -> (tf_mode, R, volume shape); TFs from scenes.random_tf"""
    from diffdvr_b200 import capi
    return {
        "preshade_identity": (capi.TF_IDENTITY, 1, (1, 6, 5, 7)),
        "preshade_texture": (capi.TF_TEXTURE, 8, (1, 9, 6, 8)),
        "preshade_linear": (capi.TF_LINEAR, 5, (1, 6, 7, 5)),
        "preshade_gaussian": (capi.TF_GAUSSIAN, 3, (1, 5, 6, 9)),
        "preshade_gaussianlog": (capi.TF_GAUSSIAN_LOG, 3, (1, 4, 5, 6)),
    }


def run_preshade(lib, case, dtype=np.float32, wrap=None, unwrap=None, sync=None):
    import scenes
    wrap = wrap or (lambda a: a)
    unwrap = unwrap or (lambda t: t)
    tf_mode, R, shape = case
    rng = np.random.RandomState(SEED + 2)
    tf = scenes.random_tf(tf_mode, R, rng, opacity=4.0).astype(dtype)[:1]
    # densities slightly outside [0, 1] too: clamps of the Identity / Linear TFs
    volume = rng.uniform(-0.1, 1.1, shape).astype(dtype)
    adj_color = rng.standard_normal((4,) + shape[1:]).astype(dtype)
    tv, tt = wrap(volume), wrap(np.ascontiguousarray(tf))
    color = wrap(np.full((4,) + shape[1:], 7.0, dtype))
    lib.preshade_volume(tv, tt, tf_mode, color)
    adj_volume = wrap(np.full(shape, 7.0, dtype))
    lib.preshade_volume_adjoint(tv, tt, tf_mode, wrap(adj_color), adj_volume)
    if sync:
        sync()
    return dict(tf=tf, volume=volume, adj_color=adj_color, color=unwrap(color), adj_volume=unwrap(adj_volume))
