"""GPU: the product's CUDA DEVICE FUNCTIONS, one at a time, against the oracles -- the function-level counterpart of the
kernel-level parity tests, on the inputs of the reference's unit tests (unittests/testAdjointBlending.cpp,
testAdjointTF.cpp, testAdjointInterpolation.cpp; tests/devfn_cases.py) plus a larger random batch.
tests/csrc/devfn_harness.cu calls blend_alpha / blend_over / blend_adjoint, stage_tf / tf_eval / tf_adjoint with both
TF-gradient accumulators, and Sampler<Trilinear | TrilinearQuad>::fetch / adjoint / flush exactly as the kernels do.
The oracles' own adjoint == finite-difference checks run in double precision in tests/test_oracle_devfn.py; here the
fp32 device functions must match the fp32 oracle: values rel <= 1e-5, adjoints rel <= 1e-4 (the tolerances of the
kernel-level suite), absolute floor 1e-6."""
import numpy as np
import pytest

import devfn_cases as dc
import oracle_loader as ol
from diffdvr_b200 import capi

pytestmark = pytest.mark.gpu
TF_MODES = {"identity": capi.TF_IDENTITY, "texture": capi.TF_TEXTURE, "linear": capi.TF_LINEAR,
            "gaussian": capi.TF_GAUSSIAN, "gaussianlog": capi.TF_GAUSSIAN_LOG}
VAL = dict(rtol=1e-5, atol=1e-6)
ADJ = dict(rtol=1e-4, atol=1e-6)


def _oracles():
    out = [dc.HostFns(ol.port(32))]
    if ol.ref(32) is not None:
        out.append(dc.HostFns(ol.ref(32)))
    return out


def _close(a, b, tol, what):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    scale = max(1.0, float(np.abs(b).max()))
    err = np.abs(a - b) - tol["rtol"] * np.abs(b)
    assert (err <= tol["atol"] * scale).all(), (what, float(err.max()))


@pytest.mark.parametrize("mode", [capi.BLEND_BEER_LAMBERT, capi.BLEND_ALPHA])
def test_blending(mode):
    dev = dc.DeviceFns()
    rng = np.random.RandomState(3)
    x = np.concatenate([dc.blend_inputs(), rng.uniform(0.01, 0.99, (2000, 9))]).astype(np.float32)
    x[100:200, 7] *= 40.0  # absorption * step > 1: the clamped branch of the Alpha blend
    got = dev.blend(mode, x)
    g = rng.uniform(-1, 1, got.shape).astype(np.float32)
    for o in _oracles():
        want = o.blend(mode, x)
        _close(got, want, VAL, "blend")
        # the adjoint consumes the ORACLE's forward output, like the kernel-level tests
        z, acc_in = dev.blend_adjoint(mode, x, want, g)
        wz, wacc = o.blend_adjoint(mode, x, want, g)
        keep = want[:, 3] < 0.999  # the inversion divides by (1 - alpha): excluded where it is ill-conditioned
        _close(z[keep], wz[keep], ADJ, "blend adjoint")
        _close(acc_in[keep], wacc[keep], dict(rtol=1e-4, atol=1e-5), "reconstructed accumulator")


def test_blending_transmittance_form():
    """blend_adjoint_T (what the Beer-Lambert adjoint kernels call: the step on T = 1 - acc.w, no quotient) against the
    oracles' blend adjoint (renderer_blending.cuh:63-98) on the same vectors; mode 100 of the harness"""
    dev = dc.DeviceFns()
    rng = np.random.RandomState(3)
    x = np.concatenate([dc.blend_inputs(), rng.uniform(0.01, 0.99, (2000, 9))]).astype(np.float32)
    for o in _oracles():
        want = o.blend(capi.BLEND_BEER_LAMBERT, x)
        g = rng.uniform(-1, 1, want.shape).astype(np.float32)
        z, acc_in = dev.blend_adjoint(100, x, want, g)
        wz, wacc = o.blend_adjoint(capi.BLEND_BEER_LAMBERT, x, want, g)
        keep = want[:, 3] < 0.999
        _close(z[keep], wz[keep], ADJ, "blend adjoint (transmittance form)")
        _close(acc_in[keep, 3], wacc[keep, 3], dict(rtol=1e-4, atol=1e-5), "reconstructed alpha")


@pytest.mark.parametrize("name", sorted(TF_MODES))
def test_transfer_functions(name):
    dev = dc.DeviceFns()
    mode = TF_MODES[name]
    rng = np.random.RandomState(5)
    for k, (table, densities) in enumerate(dc.tf_inputs(mode)):
        # beyond the unit tests' [0.01, 0.99]: the clamps at 0 and 1 and values outside
        densities = np.concatenate([densities, rng.uniform(-0.2, 1.2, 500)]).astype(np.float32)
        table = table.astype(np.float32)
        got = dev.tf(mode, table, densities)
        g = rng.uniform(-1, 1, got.shape).astype(np.float32)
        # both accumulators of the host dispatch (dvr_capi.cu): register table (R <= 6) and cached rows + atomics
        # (larger tables; exercised for every R here)
        accs = [2] + ([1] if table.shape[0] <= 6 else [])
        for o in _oracles():
            want = o.tf(mode, table, densities)
            if mode == capi.TF_GAUSSIAN_LOG:
                _close(got, want, dict(rtol=1e-5, atol=2e-5), "tf")  # log-space values of magnitude ~10
            else:
                _close(got, want, VAL, "tf")
            wd, wt = o.tf_adjoint(mode, table, densities, g)
            for acc in accs:
                ad, at = dev.tf_adjoint(mode, table, densities, g, acc)
                # piecewise-linear TFs: a density within rounding of a texel / control point may pick the other piece
                same = np.abs(ad - wd) <= 1e-4 * np.abs(wd) + 1e-4 * max(1.0, np.abs(wd).max())
                assert same.mean() >= 0.995, (name, acc, k, float(same.mean()))
                _close(at[same], wt[same], ADJ, f"adj_tf acc={acc}")


@pytest.mark.parametrize("quad", [False, True])
def test_trilinear_interpolation(quad):
    dev = dc.DeviceFns()
    rng = np.random.RandomState(9)
    volume, pos = dc.interp_inputs()
    res = np.array(volume.shape)
    # beyond the unit test: positions on and beyond every face (clamped corners), and a second, larger volume
    extra = rng.uniform(-0.6, 1.1, (400, 3)) * res
    cases = [(volume, np.concatenate([pos, extra])), (rng.uniform(0, 1, (5, 7, 6)), rng.uniform(-0.5, 7.5, (400, 3)))]
    for vol, p in cases:
        vol, p = vol.astype(np.float32), np.maximum(p, 0.0).astype(np.float32)  # the marcher never samples below 0
        if quad:
            # QUAD bakes the far-face clamps into the copy (y' / z' inside the quads, x as the duplicated slice X), so
            # its contract is what the marcher can produce: box-relative coordinates up to n - 1 plus round-off. The
            # in-place sampler keeps the reference's per-corner clamps and takes the positions a whole cell outside.
            p = np.minimum(p, (np.array(vol.shape) - 1) * (1 + 4e-7)).astype(np.float32)
        got = dev.interp(vol, p, quad)
        g = rng.uniform(-1, 1, len(p)).astype(np.float32)
        ap, av = dev.interp_adjoint(vol, p, g, quad)
        for o in _oracles():
            _close(got, o.interp(vol, p), VAL, "fetch")
            wp, wv = o.interp_adjoint(vol, p, g)
            _close(ap, wp, ADJ, "adj_pos")
            _close(av, wv, ADJ, "adj_volume")
        assert np.allclose(av.reshape(len(p), -1).sum(1), g, rtol=1e-5, atol=1e-6)  # the 8 weights sum to one


def test_quad_layout_is_bit_identical_to_in_place_sampling():
    dev = dc.DeviceFns()
    rng = np.random.RandomState(11)
    vol = rng.uniform(0, 1, (6, 5, 9)).astype(np.float32)
    # up to the far faces and the round-off beyond them (the marcher's range, see test_trilinear_interpolation)
    far = (np.array(vol.shape) - 1) * (1 + 4e-7)
    p = np.minimum(rng.uniform(0, 1.05, (2000, 3)) * np.array(vol.shape), far).astype(np.float32)
    assert (p > np.array(vol.shape) - 1).any()
    assert np.array_equal(dev.interp(vol, p, False), dev.interp(vol, p, True))
    g = rng.uniform(-1, 1, len(p)).astype(np.float32)
    (ap0, av0), (ap1, av1) = dev.interp_adjoint(vol, p, g, False), dev.interp_adjoint(vol, p, g, True)
    assert np.array_equal(ap0, ap1)
    assert np.allclose(av0, av1, rtol=1e-6, atol=1e-7)  # folded in float64 on the host
