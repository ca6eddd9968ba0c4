"""CPU: the reference's device-function unit tests run on BOTH oracles.

unittests/testAdjointBlending.cpp, testAdjointTF.cpp and testAdjointInterpolation.cpp check each adjoint against
forward differences of its forward function (unittests/check_adjoint.h) in double precision (unittests/main.cpp:22-24
warns that single precision fails). Here the same checks, with the unit tests' epsilons / precisions / margins, run on
 * the reference's own functions compiled from /root/reference (oracle/_ref/libdvr_ref64.so, oracle/ref_shim_fn.cpp),
 * the CPU port (oracle/dvr_oracle_fn.h) -- which must also agree with the reference build to rounding.
The product's CUDA device functions are compared with the fp32 oracle in tests/test_gpu_devfn.py."""
import numpy as np
import pytest

import devfn_cases as dc
import oracle_loader as ol
from diffdvr_b200 import capi

TF_MODES = {"identity": capi.TF_IDENTITY, "texture": capi.TF_TEXTURE, "linear": capi.TF_LINEAR,
            "gaussian": capi.TF_GAUSSIAN, "gaussianlog": capi.TF_GAUSSIAN_LOG}
# (epsilon, precision, margin) per test case of testAdjointTF.cpp:186-232 (no / with TF derivatives)
TF_TOL = {"identity": ((1e-6, 1e-4, 1e-6), (1e-6, 1e-4, 1e-6)), "texture": ((1e-6, 1e-4, 1e-6), (1e-6, 1e-4, 1e-6)),
          "linear": ((1e-6, 1e-4, 1e-6), (1e-6, 1e-4, 1e-6)), "gaussian": ((1e-5, 1e-3, 1e-4), (1e-5, 1e-3, 1e-3)),
          "gaussianlog": ((1e-5, 1e-3, 1e-4), (1e-6, 1e-3, 1e-4))}


def _libs():
    out = {"port": dc.HostFns(ol.port(64))}
    if ol.ref(64) is not None:
        out["reference"] = dc.HostFns(ol.ref(64))
    return out


@pytest.mark.parametrize("which", ["port", "reference"])
@pytest.mark.parametrize("mode", [capi.BLEND_BEER_LAMBERT, capi.BLEND_ALPHA])
def test_adjoint_blending(which, mode):
    """testAdjointBlending.cpp: adjoint == forward differences (1e-5, 1e-5, 1e-6) and the adjoint's reconstruction
    of the accumulator before the sample."""
    fns = _libs().get(which)
    if fns is None:
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    for x in dc.blend_inputs():
        forward = lambda v: fns.blend(mode, v[None])[0]

        def adjoint(v, e, g):
            z, acc_in = fns.blend_adjoint(mode, v[None], e[None], g[None])
            assert np.allclose(acc_in[0], v[:4], rtol=1e-10, atol=1e-12)  # REQUIRE(acc == Approx(acc_in))
            return z[0]

        dc.check_adjoint(x, forward, adjoint, 1e-5, 1e-5, 1e-6)


@pytest.mark.parametrize("which", ["port", "reference"])
@pytest.mark.parametrize("with_tf", [False, True])
@pytest.mark.parametrize("name", sorted(TF_MODES))
def test_adjoint_tf(which, name, with_tf):
    """testAdjointTF.cpp: x = (density[, tf entries]); adjoint<hasTFDerivative, false> == forward differences."""
    fns = _libs().get(which)
    if fns is None:
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    mode = TF_MODES[name]
    eps, prec, margin = TF_TOL[name][int(with_tf)]
    for table, densities in dc.tf_inputs(mode, n_tf=4, n_density=4):
        shape = table.shape
        for density in densities:
            x = np.concatenate([[density], table.ravel()]) if with_tf else np.array([density])

            def forward(v):
                t = v[1:].reshape(shape) if with_tf else table
                return fns.tf(mode, t, v[:1])[0]

            def adjoint(v, e, g):
                t = v[1:].reshape(shape) if with_tf else table
                adj_d, adj_tf = fns.tf_adjoint(mode, t, v[:1], g[None])
                return np.concatenate([adj_d, adj_tf[0].ravel()]) if with_tf else adj_d

            dc.check_adjoint(x, forward, adjoint, eps, prec, margin)


def test_tf_gaussian_log_vs_gaussian():
    """testAdjointTF.cpp:235-267: exp(TFGaussianLog) == TFGaussian"""
    for fns in _libs().values():
        for table, densities in dc.tf_inputs(capi.TF_GAUSSIAN, n_density=20):
            want = fns.tf(capi.TF_GAUSSIAN, table, densities)
            got = np.exp(fns.tf(capi.TF_GAUSSIAN_LOG, table, densities))
            assert np.allclose(got, want, rtol=1e-10, atol=1e-300)


@pytest.mark.parametrize("which", ["port", "reference"])
@pytest.mark.parametrize("with_volume", [False, True])
def test_adjoint_interpolation(which, with_volume):
    """testAdjointInterpolation.cpp: x = (position[, the 2 x 3 x 4 volume]); (1e-5, 1e-5, 1e-6)"""
    fns = _libs().get(which)
    if fns is None:
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    volume, positions = dc.interp_inputs()
    shape = volume.shape
    for pos in positions:
        x = np.concatenate([pos, volume.ravel()]) if with_volume else pos.copy()

        def forward(v):
            vol = v[3:].reshape(shape) if with_volume else volume
            return fns.interp(vol, v[None, :3])

        def adjoint(v, e, g):
            vol = v[3:].reshape(shape) if with_volume else volume
            adj_pos, adj_vol = fns.interp_adjoint(vol, v[None, :3], g)
            return np.concatenate([adj_pos[0], adj_vol[0].ravel()]) if with_volume else adj_pos[0]

        dc.check_adjoint(x, forward, adjoint, 1e-5, 1e-5, 1e-6)


@pytest.mark.parametrize("bits", [32, 64])
def test_port_functions_match_the_reference_build(bits):
    """Every function, value by value, on the unit tests' inputs: the port restates the reference to rounding."""
    ref = ol.ref(bits)
    if ref is None:
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    a, b = dc.HostFns(ol.port(bits)), dc.HostFns(ref)
    tol = dict(rtol=2e-5, atol=2e-6) if bits == 32 else dict(rtol=1e-11, atol=1e-13)
    rng = np.random.RandomState(7)
    x = dc.blend_inputs()
    for mode in (capi.BLEND_BEER_LAMBERT, capi.BLEND_ALPHA):
        e = a.blend(mode, x)
        assert np.allclose(e, b.blend(mode, x), **tol)
        g = rng.uniform(-1, 1, e.shape)
        for u, v in zip(a.blend_adjoint(mode, x, e, g), b.blend_adjoint(mode, x, e, g)):
            assert np.allclose(u, v, **tol)
    for mode in TF_MODES.values():
        for table, densities in dc.tf_inputs(mode):
            assert np.allclose(a.tf(mode, table, densities), b.tf(mode, table, densities), **tol)
            g = rng.uniform(-1, 1, (len(densities), 4))
            for u, v in zip(a.tf_adjoint(mode, table, densities, g), b.tf_adjoint(mode, table, densities, g)):
                assert np.allclose(u, v, **tol)
    volume, pos = dc.interp_inputs()
    assert np.allclose(a.interp(volume, pos), b.interp(volume, pos), **tol)
    g = rng.uniform(-1, 1, len(pos))
    for u, v in zip(a.interp_adjoint(volume, pos, g), b.interp_adjoint(volume, pos, g)):
        assert np.allclose(u, v, **tol)


def test_device_function_harness_builds_and_exports_the_same_interface():
    """tests/_build/libdvr_devfn.so (sm_100a, over the product's device headers) loads without a GPU and exports one
    devfn_* per fn_* of the oracles; no compute call here."""
    import ctypes
    import os
    import subprocess
    path = os.path.join(dc.ROOT, "tests", "_build", "libdvr_devfn.so")
    if not os.path.exists(path):
        subprocess.run(["make", "-C", os.path.join(dc.ROOT, "tests", "csrc")], check=True, capture_output=True)
    lib = ctypes.CDLL(path)
    port = ol.port(32).lib
    for fn in ("blend", "blend_adjoint", "tf", "tf_adjoint", "interp", "interp_adjoint"):
        assert hasattr(lib, "devfn_" + fn) and hasattr(port, "dvro32_fn_" + fn) and hasattr(port, "dvro64_fn_" + fn)
    sass = subprocess.run(["cuobjdump", "-lelf", path], capture_output=True, text=True).stdout
    assert "sm_100a" in sass
