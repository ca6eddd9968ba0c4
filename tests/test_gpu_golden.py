"""GPU: the CUDA path against the golden fixtures produced by the REFERENCE's own compiled host
kernels (tests/golden/ref_*.npz, see golden/make_golden.py). The reference's stepsize recurrence is
selected with compat_stepsize_overwrite=1."""
import os

import numpy as np
import pytest

import golden_cases
import gpu_runner as gr
import scenes
from diffdvr_b200 import capi

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = {name: (sc, flags, fg) for name, sc, flags, fg in golden_cases.cases()}


@pytest.mark.parametrize("name", sorted(CASES))
def test_cuda_matches_golden_reference_outputs(name):
    sc, flags, fg = CASES[name]
    gold = np.load(os.path.join(GOLDEN, f"ref_{name}.npz"))
    opts = capi.make_call_options(compat_stepsize_overwrite=True)  # the reference's stepsize recurrence, per call
    out = golden_cases.run(None, sc, flags, fg, runner=gr, forward_from=(gold["color"], gold["term"]),
                           runner_kwargs=dict(options=opts))
    import parity_log
    e_img = scenes.rel_l2(out["color"], gold["color"])
    assert e_img <= 1e-5
    # terminationIndex exact except borderline rays, < 0.1 %. The fixtures have 80..576 rays, where 0.1 %
    # is less than one ray: at most ONE ray may flip (by one sample), and the count is recorded.
    mism = out["term"] != gold["term"]
    parity_log.record("image", "strict", rel_l2=e_img, tol=1e-5, term_mismatch_rays=int(mism.sum()), rays=int(mism.size))
    assert mism.sum() <= max(1, int(1e-3 * mism.size)) and np.all(np.abs(out["term"].astype(int) - gold["term"])[mism] <= 1)
    sens = scenes.face_sensitive_rays(sc, gold["term"])
    for k in gold.files:
        if k in ("adj_volume", "adj_tf"):
            e = scenes.rel_l2(out[k], gold[k])
            parity_log.record(k, "strict", rel_l2=e, tol=1e-4)
            assert e <= 1e-4, (k, e)
        elif k.startswith("adj_"):
            # per-ray position gradients: strict on rays without a sample on a cell face
            ok = ~sens
            e_all, e_ok = scenes.rel_l2(out[k], gold[k]), scenes.rel_l2(out[k][ok], gold[k][ok])
            parity_log.record(k, "strict" if e_all <= 1e-4 else "relaxed:face-sensitive-rays", rel_l2=e_all, tol=1e-4,
                              rel_l2_unmasked=e_ok, masked_fraction=float(sens.mean()))
            assert e_ok <= 1e-4, (k, e_ok)
            assert scenes.per_ray_errors(out[k], gold[k]).max() <= 0.5
    if "fg_gradients" in gold.files:
        g, w = out["fg_gradients"], gold["fg_gradients"]
        same = out["fg_term"] == gold["fg_term"]
        ok = same & ~sens
        assert scenes.rel_l2(g[ok], w[ok]) <= 1e-4, scenes.rel_l2(g[ok], w[ok])
