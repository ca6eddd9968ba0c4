"""CPU: pins the oracle's restatement of Camera::generateRays (forward + backward) and
Renderer::compareToImage (oracle/dvr_oracle_camera.h):
  * against the functions built on the REFERENCE's own headers (oracle/_ref, ref_shim_camera.cpp),
    when that library is present (build container);
  * against the committed fixtures tests/golden/ref_camera_*.npz produced by that library;
  * against central finite differences of the forward function (fp64);
  * generated rays fed to the marcher in RayStartDir mode reproduce the ReferenceFrame render of the
    reference's compiled kernels (renderer_camera.cuh:38-58 is the same arithmetic)."""
import os

import numpy as np
import pytest

import camera_cases as cc
import oracle_loader as ol
import scenes
from diffdvr_b200 import capi

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
RAYS = cc.ray_cases()
COMPARE = cc.compare_cases()


def _close(a, b, tol):
    return scenes.rel_l2(np.asarray(a, np.float64), np.asarray(b, np.float64)) <= tol


@pytest.mark.parametrize("name", sorted(RAYS))
@pytest.mark.parametrize("bits", [32, 64])
def test_port_rays_match_reference_build(name, bits):
    ref = ol.ref(bits)
    if ref is None:
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    dt = np.float32 if bits == 32 else np.float64
    want = cc.run_rays(ref, RAYS[name], dt)
    got = cc.run_rays(ol.port(bits), RAYS[name], dt)
    # fp64: the shim's forward rays come from CameraEval (real_t tangents, renderer_camera.cuh:42-45)
    # while generateRays and the port take float tangents (camera_impl.cu:76-78): equal to fp32 round-off
    tol = 1e-6 if bits == 32 else 1e-7
    assert np.array_equal(got["ray_start"], want["ray_start"])
    assert _close(got["ray_dir"], want["ray_dir"], tol)
    assert _close(got["adj_viewport"], want["adj_viewport"], 2e-5 if bits == 32 else 1e-12)


@pytest.mark.parametrize("name", sorted(COMPARE))
@pytest.mark.parametrize("bits", [32, 64])
def test_port_compare_matches_reference_build(name, bits):
    ref = ol.ref(bits)
    if ref is None:
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    dt = np.float32 if bits == 32 else np.float64
    want = cc.run_compare(ref, COMPARE[name], dt)
    got = cc.run_compare(ol.port(bits), COMPARE[name], dt)
    for k in want:
        assert _close(got[k], want[k], 2e-6 if bits == 32 else 1e-13), k


@pytest.mark.parametrize("name", sorted(RAYS))
def test_port_rays_match_golden(name):
    gold = np.load(os.path.join(GOLDEN, f"ref_camera_{name}.npz"))
    got = cc.run_rays(ol.port(32), RAYS[name])
    assert np.array_equal(got["viewport"], gold["viewport"])  # same seeded inputs
    assert np.array_equal(got["ray_start"], gold["ray_start"])
    assert _close(got["ray_dir"], gold["ray_dir"], 1e-6)
    assert _close(got["adj_viewport"], gold["adj_viewport"], 2e-5)


@pytest.mark.parametrize("name", sorted(COMPARE))
def test_port_compare_matches_golden(name):
    gold = np.load(os.path.join(GOLDEN, f"ref_camera_{name}.npz"))
    got = cc.run_compare(ol.port(32), COMPARE[name])
    for k in gold.files:
        assert _close(got[k], gold[k], 2e-6), k


def test_rays_adjoint_matches_central_differences():
    lib = ol.port(64)
    case = dict(B=2, W=6, H=5, fovY=0.7)
    base = cc.run_rays(lib, case, np.float64)
    vp, gs, gd = base["viewport"], base["adj_ray_start"], base["adj_ray_dir"]

    def loss(v):
        s, d = np.zeros_like(gs), np.zeros_like(gd)
        lib.generate_rays(np.ascontiguousarray(v), case["fovY"], s, d)
        return (s * gs).sum(axis=(1, 2, 3)) + (d * gd).sum(axis=(1, 2, 3))  # per view

    fd = np.zeros_like(vp)
    eps = 1e-6
    for r in range(3):
        for c in range(3):
            hi, lo = vp.copy(), vp.copy()
            hi[:, r, c] += eps
            lo[:, r, c] -= eps
            fd[:, r, c] = (loss(hi) - loss(lo)) / (2 * eps)
    assert scenes.rel_l2(base["adj_viewport"], fd) <= 1e-7


def test_compare_gradient_is_directional_derivative():
    lib = ol.port(64)
    case = dict(B=1, W=5, H=4, D=3)
    out = cc.run_compare(lib, case, np.float64)
    rng = np.random.RandomState(cc.SEED + 1)
    color = rng.uniform(0, 1, (1, 4, 5, 4))
    reference = rng.uniform(0, 1, (1, 4, 5, 4))
    grads = rng.standard_normal((1, 4, 5, 3, 4))
    eps = 1e-6
    for d in range(3):
        hi = np.linalg.norm(color + eps * grads[..., d, :] - reference, axis=-1)
        lo = np.linalg.norm(color - eps * grads[..., d, :] - reference, axis=-1)
        assert np.allclose(out["gradients"][..., d], (hi - lo) / (2 * eps), rtol=1e-6, atol=1e-8)
    assert np.allclose(out["difference"][..., 0], np.linalg.norm(color - reference, axis=-1))
    assert np.allclose(out["difference_reduced"][0], out["difference"].mean())
    assert np.allclose(out["gradients_reduced"], out["gradients"].mean(axis=(0, 1, 2)))


def test_generated_rays_reproduce_reference_frame_render():
    """RayStartDir render of generate_rays output == ReferenceFrame render (reference kernels if built)."""
    lib = ol.ref(32) or ol.port(32)
    sc = scenes.make_scene(n=16, W=12, H=10, B=2, camera_mode=capi.CAMERA_REFERENCE_FRAME, tf_mode=capi.TF_TEXTURE,
                           R=8, seed=5, opacity=4.0)
    color_frame, term_frame = scenes.np_forward(lib, sc)
    start = np.zeros((sc.B, sc.H, sc.W, 3), np.float32)
    direction = np.zeros_like(start)
    ol.port(32).generate_rays(np.ascontiguousarray(sc.camera_matrix), sc.fov_y, start, direction)
    sc.camera_mode = capi.CAMERA_RAY_START_DIR
    sc.ray_start, sc.ray_dir = start, direction
    color_rays, term_rays = scenes.np_forward(lib, sc)
    assert scenes.rel_l2(color_rays, color_frame) <= 1e-6
    assert (term_rays != term_frame).mean() <= 0.01
