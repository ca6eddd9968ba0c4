"""GPU: the two volume layouts of the trilinear sampler -- the caller's strided tensor sampled in place
("volume_layout" = 0) and the library's QUAD copy (= 1; two 16-byte loads per cell) -- give the same results:
the forward image bit-for-bit (same corner values, same arithmetic), the gradients up to atomic summation order;
both against the golden fixtures of the reference's compiled kernels."""
import os

import numpy as np
import pytest

import golden_cases
import gpu_runner as gr
import scenes
from diffdvr_b200 import capi

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = {name: (sc, flags, fg) for name, sc, flags, fg in golden_cases.cases()
         if sc.filter_mode == capi.FILTER_TRILINEAR}


def _run(layout, sc, flags, gold):
    lib = capi.load()
    opts = capi.make_call_options(volume_layout=layout, compat_stepsize_overwrite=True)
    n0 = lib.launch_count()
    color, term = gr.forward(sc, lib, options=opts)
    launches = lib.launch_count() - n0
    adj_color = np.random.RandomState(golden_cases.ADJ_SEED).randn(*color.shape).astype(np.float32)
    adj = gr.adjoint(sc, gold["color"], gold["term"], adj_color, lib=lib, options=opts, **flags)
    return color, term, adj, launches


@pytest.mark.parametrize("name", sorted(CASES))
def test_quad_copy_equals_in_place_sampling(name):
    sc, flags, _ = CASES[name]
    gold = np.load(os.path.join(GOLDEN, f"ref_{name}.npz"))
    c0, t0, a0, l0 = _run(0, sc, flags, gold)
    c1, t1, a1, l1 = _run(1, sc, flags, gold)
    assert l0 == 1 and l1 == 2  # the copy is one extra launch
    assert np.array_equal(c0, c1) and np.array_equal(t0, t1)
    assert scenes.rel_l2(c1, gold["color"]) <= 1e-5
    for k in a0:
        assert scenes.rel_l2(a1[k], a0[k]) <= 2e-6, k
    for k in ("adj_volume", "adj_tf"):
        if k in a1:
            assert scenes.rel_l2(a1[k], gold[k]) <= 1e-4, k


def test_automatic_choice_follows_the_work_estimate():
    lib = capi.load()
    big = scenes.make_scene(n=48, W=4, H=4, B=1, seed=1)      # 16 rays, 110592 voxels: in place
    small = scenes.make_scene(n=16, W=32, H=32, B=1, seed=1)  # 1024 rays, 4096 voxels: copy
    for sc, want in ((big, 1), (small, 2)):
        n0 = lib.launch_count()
        gr.forward(sc, lib)
        assert lib.launch_count() - n0 == want


@pytest.mark.parametrize("name", sorted(n for n, (_, _, fg) in CASES.items() if fg is not None))
def test_forward_mode_quad_copy_equals_in_place_sampling(name):
    """Forward-mode differentiation samples through the same two layouts: identical images and termination indices,
    Jacobians equal to rounding (the dual-number x-lerp is evaluated from the same corner values) and equal to the
    reference's fixtures within 1e-4."""
    sc, _, fg = CASES[name]
    st = golden_cases.fg_settings(sc, fg)
    gold = np.load(os.path.join(GOLDEN, f"ref_{name}.npz"))
    lib = capi.load()
    res = {}
    for layout in (0, 1):
        n0 = lib.launch_count()
        res[layout] = gr.forward_gradients(sc, lib=lib, options=capi.make_call_options(volume_layout=layout), **st)
        launches = lib.launch_count() - n0
        if st.get("d_tf") is None:
            passes = 1                                  # ray / stepsize tangents: the input-major pass
        else:
            passes = 2                                  # colour (plain forward or input-major) + the sample-major Jacobian
        assert launches == passes + layout  # ONE copy serves every pass
    (c0, t0, g0), (c1, t1, g1) = res[0], res[1]
    assert np.array_equal(c0, c1) and np.array_equal(t0, t1)
    assert scenes.rel_l2(g1, g0) <= 2e-6
    same = t1 == gold["fg_term"]
    assert scenes.rel_l2(g1[same], gold["fg_gradients"][same]) <= 1e-4
