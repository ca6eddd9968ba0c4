"""CPU: pins the oracle's restatement of TFUtils::preshadeVolume and its density adjoint
(oracle/dvr_oracle_tfutils.h) against the reference-built shim (oracle/_ref: the reference's own
TransferFunctionEval<mode>::eval / adjoint), the committed fixtures tests/golden/ref_preshade_*.npz, and central
finite differences of the forward function."""
import os

import numpy as np
import pytest

import camera_cases as cc
import oracle_loader as ol
import scenes
from diffdvr_b200 import capi

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = cc.preshade_cases()


@pytest.mark.parametrize("name", sorted(CASES))
@pytest.mark.parametrize("bits", [32, 64])
def test_port_matches_reference_build(name, bits):
    ref = ol.ref(bits)
    if ref is None:
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    dt = np.float32 if bits == 32 else np.float64
    want = cc.run_preshade(ref, CASES[name], dt)
    got = cc.run_preshade(ol.port(bits), CASES[name], dt)
    tol = 2e-6 if bits == 32 else 1e-12
    assert scenes.rel_l2(got["color"], want["color"]) <= tol
    assert scenes.rel_l2(got["adj_volume"], want["adj_volume"]) <= 10 * tol


@pytest.mark.parametrize("name", sorted(CASES))
def test_port_matches_golden(name):
    gold = np.load(os.path.join(GOLDEN, f"ref_{name}.npz"))
    got = cc.run_preshade(ol.port(32), CASES[name])
    assert np.array_equal(got["tf"], gold["tf"])
    assert scenes.rel_l2(got["color"], gold["color"]) <= 2e-6
    assert scenes.rel_l2(got["adj_volume"], gold["adj_volume"]) <= 2e-5


@pytest.mark.parametrize("name", ["preshade_texture", "preshade_gaussian", "preshade_gaussianlog"])
def test_adjoint_matches_central_differences(name):
    lib = ol.port(64)
    tf_mode, _, shape = CASES[name]
    out = cc.run_preshade(lib, CASES[name], np.float64)
    vol, tf, adj = out["volume"], np.ascontiguousarray(out["tf"]), out["adj_color"]

    def colour(v):
        c = np.zeros((4,) + shape[1:])
        lib.preshade_volume(np.ascontiguousarray(v), tf, tf_mode, c)
        return c

    eps = 1e-6
    fd = ((colour(vol + eps) - colour(vol - eps)) / (2 * eps) * adj).sum(axis=0, keepdims=True)
    # the texture TF is piecewise linear: skip voxels whose +-eps stencil straddles a texel boundary
    ok = np.ones(shape, bool)
    if tf_mode == capi.TF_TEXTURE:
        R = tf.shape[1]
        d = vol * R - 0.5
        ok = np.abs(d - np.round(d)) > 1e-4
    assert scenes.rel_l2(out["adj_volume"][ok], fd[ok]) <= 1e-6


def test_texture_preshade_equals_numpy_lookup():
    """Texture TF: every pre-shaded voxel is the linear interpolation of the two texels around density*R - 0.5
    (renderer_tf.cuh:160-171), written out in numpy."""
    rng = np.random.RandomState(3)
    R = 6
    tf = scenes.random_tf(capi.TF_TEXTURE, R, rng, opacity=4.0)
    volume = rng.uniform(-0.2, 1.2, (1, 7, 6, 5)).astype(np.float32)
    color = np.zeros((4,) + volume.shape[1:], np.float32)
    ol.port(32).preshade_volume(volume, np.ascontiguousarray(tf), capi.TF_TEXTURE, color)
    d = volume[0] * R - 0.5
    fl = np.floor(d)
    i0 = np.clip(fl.astype(int), 0, R - 1)
    i1 = np.clip(fl.astype(int) + 1, 0, R - 1)
    f = (d - fl)[..., None]
    want = ((1 - f) * tf[0][i0] + f * tf[0][i1]).transpose(3, 0, 1, 2)
    assert np.allclose(color, want, atol=1e-5)
