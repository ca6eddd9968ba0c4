"""CPU tests (no GPU): the oracle (oracle/dvr_oracle.cpp, the CPU restatement) is PINNED against
  1. the reference's own known-answer table unittests/NonConvexDensity.tsv (golden/NonConvexDensity.tsv);
  2. golden fixtures produced by the reference's own compiled host kernels (golden/ref_*.npz,
     generator golden/make_golden.py);
  3. the reference's compiled host kernels directly, when oracle/_ref is present (all mode combinations);
  4. the reference's own consistency property forward-mode == adjoint (unittests/testConsistency*.cpp);
  5. central finite differences in fp64 (the reference's check_adjoint.h uses one-sided differences).
"""
import os

import numpy as np
import pytest

import golden_cases
import oracle_loader as ol
import scenes
from diffdvr_b200 import capi

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.fixture()
def reference_stepsize_semantics():
    ol.port_set_option("compat_stepsize_overwrite", 1)
    yield
    ol.port_set_option("compat_stepsize_overwrite", 0)


@pytest.fixture()
def complete_stepsize_semantics():
    ol.port_set_option("compat_stepsize_overwrite", 0)
    yield


# ---- 1. NonConvexDensity.tsv ---------------------------------------------------------------------
def read_tsv():
    rows = []
    with open(os.path.join(GOLDEN, "NonConvexDensity.tsv")) as f:
        for line in f.read().splitlines()[2:]:
            if line.strip():
                rows.append([float(v) for v in line.split("\t")])
    return np.array(rows)


@pytest.mark.parametrize("which", ["port", "ref"])
def test_nonconvex_density_golden_table(which):
    """d1, red, alpha, loss, dloss/dd1 printed with 5 decimals (unittests/testNonConvexDensity.cpp:60-78).
    Pins the Gaussian TF + Beer-Lambert blend + one-tangent dual arithmetic."""
    lib = ol.port(32) if which == "port" else ol.ref(32)
    if lib is None:
        pytest.skip("oracle/_ref not built")
    table = read_tsv()
    assert table.shape == (130, 5)
    got = np.array([lib.nonconvex_density(d1) for d1 in table[:, 0]])
    # 5 printed decimals -> half a unit in the last place, plus fp32 round-off of the 33..1000-step sums
    assert np.max(np.abs(got[:, 0] - table[:, 1])) <= 6e-6 + 1e-6
    assert np.max(np.abs(got[:, 1] - table[:, 2])) <= 6e-6 + 1e-6
    assert np.max(np.abs(got[:, 2] - table[:, 3])) <= 6e-6 + 2e-6
    # d1 = d0 = -1: loss = |0| and its derivative is 0/0 (the reference prints -nan(ind) / nan there)
    ok = np.isfinite(table[:, 4]) & np.isfinite(got[:, 3])
    assert ok.sum() >= 128
    assert np.max(np.abs(got[ok, 3] - table[ok, 4])) <= 6e-6 + 5e-6


# ---- 2. golden fixtures of the reference's compiled kernels ------------------------------------------
CASES = {name: (sc, flags, fg) for name, sc, flags, fg in golden_cases.cases()}


def compare_with_golden(name, out, gold, img_tol=1e-6, grad_tol=2e-5):
    assert scenes.rel_l2(out["color"], gold["color"]) <= img_tol, name
    assert np.array_equal(out["term"], gold["term"]), name
    for k in gold.files:
        if k.startswith("adj_"):
            assert scenes.rel_l2(out[k], gold[k]) <= grad_tol, (name, k, scenes.rel_l2(out[k], gold[k]))
    if "fg_gradients" in gold.files:
        assert np.array_equal(out["fg_term"], gold["fg_term"])
        assert scenes.rel_l2(out["fg_color"], gold["fg_color"]) <= img_tol
        assert scenes.rel_l2(out["fg_gradients"], gold["fg_gradients"]) <= grad_tol, (
            name, scenes.rel_l2(out["fg_gradients"], gold["fg_gradients"]))


@pytest.mark.parametrize("name", sorted(CASES))
def test_port_matches_golden_reference_outputs(name, reference_stepsize_semantics):
    sc, flags, fg = CASES[name]
    gold = np.load(os.path.join(GOLDEN, f"ref_{name}.npz"))
    out = golden_cases.run(ol.port(32), sc, flags, fg)
    compare_with_golden(name, out, gold)


# ---- 3. the reference's compiled host kernels, every mode combination ---------------------------------
MODES = [(f, c, t, b)
         for f in (capi.FILTER_NEAREST, capi.FILTER_TRILINEAR)
         for c in (capi.CAMERA_RAY_START_DIR, capi.CAMERA_INVERSE_VIEW_MATRIX, capi.CAMERA_REFERENCE_FRAME)
         for t in (capi.TF_IDENTITY, capi.TF_TEXTURE, capi.TF_LINEAR, capi.TF_GAUSSIAN, capi.TF_GAUSSIAN_LOG)
         for b in (capi.BLEND_BEER_LAMBERT, capi.BLEND_ALPHA)]


@pytest.mark.parametrize("filt,cam,tfm,blend", MODES)
def test_port_matches_reference_host_kernels(filt, cam, tfm, blend, reference_stepsize_semantics):
    ref = ol.ref(32)
    if ref is None:
        pytest.skip("oracle/_ref not built (needs /root/reference at build time)")
    sc = scenes.make_scene(n=14, W=11, H=9, B=2, camera_mode=cam, tf_mode=tfm, R=4, blend_mode=blend,
                           filter_mode=filt, seed=31, opacity=3.0 if tfm != capi.TF_GAUSSIAN_LOG else 1.0)
    flags = dict(volume=True, step=True, camera=True, tf=True)
    fg = None if tfm == capi.TF_GAUSSIAN_LOG else "tf"   # forward-mode is deleted for GaussianLog (renderer_tf.cuh:692-702)
    want = golden_cases.run(ref, sc, flags, fg)
    got = golden_cases.run(ol.port(32), sc, flags, fg)
    assert np.array_equal(got["term"], want["term"])
    assert scenes.rel_l2(got["color"], want["color"]) <= 1e-6
    for k in want:
        if k.startswith("adj_") or k == "fg_gradients":
            assert scenes.rel_l2(got[k], want[k]) <= 2e-5, (k, scenes.rel_l2(got[k], want[k]))


def test_fp64_port_matches_fp64_reference_build(reference_stepsize_semantics):
    ref = ol.ref(64)
    if ref is None:
        pytest.skip("oracle/_ref not built")
    sc = scenes.make_scene(n=14, W=11, H=9, B=1, tf_mode=capi.TF_LINEAR, R=4, seed=32, opacity=3.0)
    c_r, t_r = scenes.np_forward(ref, sc, np.float64)
    c_p, t_p = scenes.np_forward(ol.port(64), sc, np.float64)
    assert np.array_equal(t_r, t_p) and scenes.rel_l2(c_p, c_r) <= 1e-13
    adj_color = np.random.RandomState(1).randn(*c_r.shape)
    kw = dict(volume=True, step=True, camera=True, tf=True, dtype=np.float64)
    a_r = scenes.np_adjoint(ref, sc, c_r, t_r, adj_color, **kw)
    a_p = scenes.np_adjoint(ol.port(64), sc, c_r, t_r, adj_color, **kw)
    for k in a_r:
        assert scenes.rel_l2(a_p[k], a_r[k]) <= 1e-11, k


# ---- 4. forward-mode == adjoint (the reference's consistency tests), fp64 --------------------------------
def fwd_to_adjoint(lib, sc, D, st, adj_color, dtype):
    B, H, W = sc.B, sc.H, sc.W
    inp = sc.inputs(lambda a: a.astype(dtype) if a.dtype != np.int32 else a)
    s = capi.make_fd_settings(D, d_tf=st.get("d_tf"), has_tf_derivatives="d_tf" in st,
                              **{k: v for k, v in st.items() if k != "d_tf"})
    color, term = np.zeros((B, H, W, 4), dtype), np.zeros((B, H, W), np.int32)
    g = np.zeros((B, H, W, D, 4), dtype)
    lib.render_forward_gradients(inp, s, color, term, g)
    out = {}
    if st.get("d_stepsize", -1) >= 0:
        out["adj_step_size"] = np.zeros((B, H, W), dtype)
    if "d_ray_start" in st:
        out["adj_camera_ray_start"] = np.zeros((B, H, W, 3), dtype)
        out["adj_camera_ray_dir"] = np.zeros((B, H, W, 3), dtype)
    if "d_tf" in st:
        out["adj_tf"] = np.zeros(sc.tf.shape, dtype)
    lib.forward_variables_to_gradients(g, adj_color.astype(dtype), s, capi.make_adjoint_outputs(**out))
    return color, term, out


@pytest.mark.parametrize("tfm,R", [(capi.TF_IDENTITY, 1), (capi.TF_TEXTURE, 2), (capi.TF_TEXTURE, 4),
                                    (capi.TF_TEXTURE, 8), (capi.TF_LINEAR, 4), (capi.TF_GAUSSIAN, 3)])
def test_forward_mode_equals_adjoint_fp64(tfm, R, complete_stepsize_semantics):
    """unittests/testConsistencyTF.cpp:204-222, testConsistencyCamera.cpp:205-217 (EPS = 1e-5 symmetric
    error); with the complete stepsize recurrence the stepsize gradient agrees as well."""
    o64 = ol.port(64)
    for sc, st in ((scenes.consistency_tf_scene(tfm, R, seed=R, W=32, H=32), "tf"),
                   (scenes.consistency_camera_scene(tfm, R, seed=R),
                    dict(D=7, d_ray_start=(0, 1, 2), d_ray_dir=(3, 4, 5), d_stepsize=6))):
        st = golden_cases.fg_settings(sc, st)
        D = st.pop("D")
        adj_color = np.random.RandomState(5).randn(sc.B, sc.H, sc.W, 4)
        color, term, fwd = fwd_to_adjoint(o64, sc, D, st, adj_color, np.float64)
        adj = scenes.np_adjoint(o64, sc, color, term, adj_color, dtype=np.float64, tf="d_tf" in st,
                                camera="d_ray_start" in st, step="d_stepsize" in st)
        for k in fwd:
            assert scenes.symmetric_error(fwd[k], adj[k]) <= 1e-5, (k, scenes.symmetric_error(fwd[k], adj[k]))


# ---- 5. central finite differences, fp64 -----------------------------------------------------------------
def loss_of(lib, sc, adj_color):
    c, _ = scenes.np_forward(lib, sc, np.float64)
    return float((c * adj_color).sum())


@pytest.mark.parametrize("tfm,blend", [(capi.TF_TEXTURE, capi.BLEND_BEER_LAMBERT), (capi.TF_LINEAR, capi.BLEND_ALPHA),
                                       (capi.TF_GAUSSIAN, capi.BLEND_BEER_LAMBERT),
                                       (capi.TF_IDENTITY, capi.BLEND_BEER_LAMBERT)])
def test_adjoint_against_central_differences(tfm, blend, complete_stepsize_semantics):
    """d<adj_color, color>/d{volume, tf, stepsize, rayStart, rayDir} vs (f(x+h)-f(x-h))/2h in fp64.
    early_out = 2 disables the early ray termination, whose discrete exit is not differentiable."""
    o64 = ol.port(64)
    sc = scenes.make_scene(n=8, W=6, H=5, B=1, camera_mode=capi.CAMERA_RAY_START_DIR, tf_mode=tfm, R=4,
                           blend_mode=blend, seed=41, opacity=2.0, volume_kind="rand", step_voxels=0.37)
    sc.early_out = 2.0
    sc.volume = (0.1 + 0.8 * sc.volume).astype(np.float32)
    sc.step_size = np.full((1, sc.H, sc.W), sc.step_size, dtype=np.float32)
    rng = np.random.RandomState(3)
    # rays start INSIDE the box (tmin = 0): the reference does not differentiate the box intersection
    # (renderer_kernels.cuh:201-208 uses values only), so entry points must not move with the ray
    sc.ray_start = (-0.35 + 0.3 * rng.rand(1, sc.H, sc.W, 3)).astype(np.float32)
    d = rng.randn(1, sc.H, sc.W, 3)
    sc.ray_dir = (d / np.linalg.norm(d, axis=-1, keepdims=True)).astype(np.float32)
    adj_color = rng.randn(1, sc.H, sc.W, 4)
    color, term = scenes.np_forward(o64, sc, np.float64)
    assert term.min() > 3
    adj = scenes.np_adjoint(o64, sc, color, term, adj_color, volume=True, step=True, camera=True, tf=True,
                            dtype=np.float64)
    base_term = term.copy()

    def fd(array, index, h):
        # perturb in fp64: Scene converts with astype(float64), so hold the array as float64
        old = array[index]
        array[index] = old + h
        cp, tp = scenes.np_forward(o64, sc, np.float64)
        array[index] = old - h
        cm, tm = scenes.np_forward(o64, sc, np.float64)
        array[index] = old
        if not (np.array_equal(tp, base_term) and np.array_equal(tm, base_term)):
            return None  # a sample crossed the exit face: not differentiable there
        return float(((cp - cm) * adj_color).sum()) / (2 * h)

    def check(name, array, grad, picks, h, rtol):
        n_ok = 0
        for index in picks:
            want = fd(array, index, h)
            if want is None:
                continue
            got = grad[index]
            assert abs(got - want) <= rtol * max(1.0, abs(want)) , (name, index, got, want)
            n_ok += 1
        assert n_ok >= len(picks) // 2, name

    sc.volume = sc.volume.astype(np.float64)
    sc.tf = sc.tf.astype(np.float64)
    sc.step_size = sc.step_size.astype(np.float64)
    sc.ray_start = sc.ray_start.astype(np.float64)
    sc.ray_dir = sc.ray_dir.astype(np.float64)
    vol_picks = [tuple([0] + list(rng.randint(1, 7, size=3))) for _ in range(12)]
    check("adj_volume", sc.volume, adj["adj_volume"], vol_picks, 1e-6, 1e-5)
    tf_picks = [(0, r, c) for r in range(sc.tf.shape[1]) for c in range(sc.tf.shape[2])]
    if tfm == capi.TF_LINEAR:  # moving the end positions changes the search interval
        tf_picks = [p for p in tf_picks if not (p[2] == 4 and p[1] in (0, sc.tf.shape[1] - 1))]
    check("adj_tf", sc.tf, adj["adj_tf"], tf_picks, 1e-6, 2e-5)
    ray_picks = [(0, int(rng.randint(sc.H)), int(rng.randint(sc.W))) for _ in range(6)]
    check("adj_step_size", sc.step_size, adj["adj_step_size"], ray_picks, 1e-7, 1e-4)
    check("adj_camera_ray_start", sc.ray_start, adj["adj_camera_ray_start"],
          [p + (c,) for p in ray_picks for c in range(3)], 1e-7, 1e-4)
    check("adj_camera_ray_dir", sc.ray_dir, adj["adj_camera_ray_dir"],
          [p + (c,) for p in ray_picks for c in range(3)], 1e-7, 1e-4)


def test_reference_stepsize_recurrence_differs_from_complete_gradient():
    """the reference overwrites the blend term of adj_stepSize every sample; the port
    reproduces it (compat=1, pinned by the golden fixtures) and offers the complete recurrence."""
    sc = scenes.make_scene(n=12, W=8, H=6, B=1, tf_mode=capi.TF_TEXTURE, R=4, seed=2, opacity=3.0)
    o32 = ol.port(32)
    c, t = scenes.np_forward(o32, sc)
    adj_color = np.random.RandomState(2).randn(*c.shape).astype(np.float32)
    try:
        ol.port_set_option("compat_stepsize_overwrite", 1)
        a = scenes.np_adjoint(o32, sc, c, t, adj_color, step=True)["adj_step_size"]
        ol.port_set_option("compat_stepsize_overwrite", 0)
        b = scenes.np_adjoint(o32, sc, c, t, adj_color, step=True)["adj_step_size"]
    finally:
        ol.port_set_option("compat_stepsize_overwrite", 0)
    assert scenes.rel_l2(a, b) > 1e-2


def test_empty_and_degenerate_inputs():
    """rays that miss the box render (0,0,0,0) with terminationIndex 0; 1x1 images; R=1 texture."""
    o32 = ol.port(32)
    sc = scenes.make_scene(n=8, W=1, H=1, B=1, tf_mode=capi.TF_TEXTURE, R=1, seed=1)
    c, t = scenes.np_forward(o32, sc)
    assert c.shape == (1, 1, 1, 4) and t[0, 0, 0] > 0
    sc = scenes.make_scene(n=8, W=5, H=4, B=1, camera_mode=capi.CAMERA_RAY_START_DIR, seed=1)
    sc.ray_dir = -sc.ray_dir  # looking away from the box
    c, t = scenes.np_forward(o32, sc)
    assert not c.any() and not t.any()
    a = scenes.np_adjoint(o32, sc, c, t, np.ones_like(c), volume=True, tf=True, camera=True, step=True)
    assert all(not v.any() for v in a.values())
