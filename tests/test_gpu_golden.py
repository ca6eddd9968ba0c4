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
    lib = capi.load()
    lib.set_option("compat_stepsize_overwrite", 1)
    try:
        out = golden_cases.run(None, sc, flags, fg, runner=gr, forward_from=(gold["color"], gold["term"]))
    finally:
        lib.set_option("compat_stepsize_overwrite", 0)
    assert scenes.rel_l2(out["color"], gold["color"]) <= 1e-5
    mism = out["term"] != gold["term"]
    assert mism.mean() <= 0.01 and np.all(np.abs(out["term"].astype(int) - gold["term"])[mism] <= 1)
    sens = scenes.face_sensitive_rays(sc, gold["term"])
    for k in gold.files:
        if k in ("adj_volume", "adj_tf"):
            assert scenes.rel_l2(out[k], gold[k]) <= 1e-4, (k, scenes.rel_l2(out[k], gold[k]))
        elif k.startswith("adj_"):
            # per-ray position gradients: strict on rays without a sample on a cell face
            ok = ~sens
            assert scenes.rel_l2(out[k][ok], gold[k][ok]) <= 1e-4, (k, scenes.rel_l2(out[k][ok], gold[k][ok]))
            assert scenes.per_ray_errors(out[k], gold[k]).max() <= 0.5
    if "fg_gradients" in gold.files:
        g, w = out["fg_gradients"], gold["fg_gradients"]
        same = out["fg_term"] == gold["fg_term"]
        ok = same & ~sens
        assert scenes.rel_l2(g[ok], w[ok]) <= 1e-4, scenes.rel_l2(g[ok], w[ok])
