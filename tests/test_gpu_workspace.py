"""GPU: per-call options and the volume-layout workspace of ABI version 2 (include/dvr_b200.h DvrCallOptions /
DvrWorkspace): a cached QUAD copy gives the bits of a per-call copy, is rebuilt exactly when the caller says the volume
changed, one accumulation scope equals the sum of per-call adjoints, and every fallback to the in-place path is
counted and still correct."""
import numpy as np
import pytest
import torch

import gpu_runner as gr
import scenes
from diffdvr_b200 import capi

pytestmark = pytest.mark.gpu


def scene(**kw):
    args = dict(n=24, W=40, H=32, B=2, tf_mode=capi.TF_TEXTURE, R=8, seed=3, opacity=3.0)
    args.update(kw)
    return scenes.make_scene(**args)


def render(lib, inp, B, H, W, options=None):
    color = torch.empty(B, H, W, 4, device="cuda")
    term = torch.empty(B, H, W, dtype=torch.int32, device="cuda")
    lib.render_forward(inp, color, term, options=options)
    torch.cuda.synchronize()
    return color, term


def test_cached_copy_is_reused_and_rebuilt_on_demand():
    lib = capi.load()
    sc = scene()
    inp = sc.inputs(gr.to_cuda)
    ws = lib.workspace()
    opts = capi.make_call_options(volume_layout=1, workspace=ws)
    ref_c, ref_t = render(lib, inp, sc.B, sc.H, sc.W, capi.make_call_options(volume_layout=1))
    b0, r0 = lib.stat("quad_builds"), lib.stat("quad_reuses")
    c1, t1 = render(lib, inp, sc.B, sc.H, sc.W, opts)      # builds into the workspace
    c2, t2 = render(lib, inp, sc.B, sc.H, sc.W, opts)      # served by the cached copy: one launch
    assert lib.stat("quad_builds") - b0 == 1 and lib.stat("quad_reuses") - r0 == 1
    assert torch.equal(c1, ref_c) and torch.equal(c2, ref_c) and torch.equal(t2, ref_t)
    n = inp.volume.shape[1]
    assert ws.bytes() == 16 * (n + 1) * n * n       # X + 1 slices: the far-face clamp of x is a duplicated slice
    # contents change under the same pointer: stale until the caller bumps the "version"
    inp.volume.mul_(0.5)
    fresh_c, _ = render(lib, inp, sc.B, sc.H, sc.W, capi.make_call_options(volume_layout=1))
    stale_c, _ = render(lib, inp, sc.B, sc.H, sc.W, opts)
    assert torch.equal(stale_c, ref_c) and not torch.equal(fresh_c, ref_c)
    ws.set_volume(inp.volume)
    c3, _ = render(lib, inp, sc.B, sc.H, sc.W, opts)
    assert torch.equal(c3, fresh_c)
    ws.invalidate()
    b1 = lib.stat("quad_builds")
    c4, _ = render(lib, inp, sc.B, sc.H, sc.W, opts)
    assert lib.stat("quad_builds") - b1 == 1 and torch.equal(c4, fresh_c)
    # a different tensor (other pointer) never hits the cache
    other = inp.volume.clone()
    inp.volume = other
    c5, _ = render(lib, inp, sc.B, sc.H, sc.W, opts)
    assert torch.equal(c5, fresh_c) and lib.stat("quad_builds") - b1 == 2
    # automatic layout: a valid cached copy is always used, even for a call too small to pay for a build
    tiny = scene(W=4, H=4, B=1)
    tin = tiny.inputs(gr.to_cuda)
    tin.volume = other
    n0 = lib.launch_count()
    render(lib, tin, 1, 4, 4, capi.make_call_options(workspace=ws))
    assert lib.launch_count() - n0 == 1
    ws.close()


def test_one_accumulation_scope_equals_the_sum_of_per_call_adjoints():
    lib = capi.load()
    sc = scene()
    inp = sc.inputs(gr.to_cuda)
    color, term = render(lib, inp, sc.B, sc.H, sc.W)
    g = torch.Generator(device="cuda").manual_seed(1)
    adj_colors = [torch.randn(color.shape, device="cuda", generator=g) for _ in range(3)]
    want = torch.zeros_like(inp.volume)
    for a in adj_colors:
        lib.render_adjoint(inp, color, term, a, capi.make_adjoint_outputs(adj_volume=want),
                           options=capi.make_call_options(volume_layout=1))
    ws = lib.workspace()
    opts = capi.make_call_options(volume_layout=1, workspace=ws)
    got = torch.zeros_like(inp.volume)
    f0, n0 = lib.stat("folds"), lib.launch_count()
    ws.begin_volume_gradient(got)
    for a in adj_colors:
        lib.render_adjoint(inp, color, term, a, capi.make_adjoint_outputs(adj_volume=got), options=opts)
    assert float(got.abs().max()) == 0.0                     # nothing reaches adj_volume before the fold
    ws.end_volume_gradient(got)
    torch.cuda.synchronize()
    assert lib.stat("folds") - f0 == 1
    assert lib.launch_count() - n0 == 1 + 3 + 1             # one copy, three marches, one fold
    assert float((got - want).norm() / want.norm()) <= 2e-6
    # outside a scope the workspace's scratch is zeroed and folded per call
    again = torch.zeros_like(inp.volume)
    for a in adj_colors:
        lib.render_adjoint(inp, color, term, a, capi.make_adjoint_outputs(adj_volume=again), options=opts)
    torch.cuda.synchronize()
    assert lib.stat("folds") - f0 == 4 and float((again - want).norm() / want.norm()) <= 2e-6
    # scope misuse is an error, not a silent wrong gradient
    ws.begin_volume_gradient(got)
    with pytest.raises(capi.DvrError, match="already open"):
        ws.begin_volume_gradient(got)
    with pytest.raises(capi.DvrError, match="differs from the tensor"):
        ws.end_volume_gradient(again)
    ws.end_volume_gradient(got)
    with pytest.raises(capi.DvrError, match="no volume-gradient scope"):
        ws.end_volume_gradient(got)
    ws.close()


def test_fallbacks_to_the_in_place_path_are_counted_and_correct():
    lib = capi.load()
    sc = scene(B=1)
    inp = sc.inputs(gr.to_cuda)
    color, term = render(lib, inp, 1, sc.H, sc.W)
    adj_color = torch.randn(color.shape, device="cuda", generator=torch.Generator(device="cuda").manual_seed(2))
    want = torch.zeros_like(inp.volume)
    lib.render_adjoint(inp, color, term, adj_color, capi.make_adjoint_outputs(adj_volume=want),
                       options=capi.make_call_options(volume_layout=0))
    # scratch above the per-call cap
    f0 = lib.stat("quad_fallbacks")
    capped = torch.zeros_like(inp.volume)
    opts = capi.make_call_options(volume_layout=1, max_scratch_bytes=1024)
    c2, _ = render(lib, inp, 1, sc.H, sc.W, opts)
    lib.render_adjoint(inp, color, term, adj_color, capi.make_adjoint_outputs(adj_volume=capped), options=opts)
    torch.cuda.synchronize()
    assert lib.stat("quad_fallbacks") - f0 == 2 and torch.equal(c2, color)
    assert float((capped - want).norm() / want.norm()) <= 2e-6
    # adj_volume with a stride-0 spatial dimension (an expanded view): every voxel column shares one address, which
    # only the atomics of the in-place path accumulate correctly
    n = inp.volume.shape[1]
    base = torch.zeros(1, n, n, 1, device="cuda")
    f1 = lib.stat("quad_fallbacks")
    lib.render_adjoint(inp, color, term, adj_color, capi.make_adjoint_outputs(adj_volume=base.expand(1, n, n, n)),
                       options=capi.make_call_options(volume_layout=1))
    torch.cuda.synchronize()
    assert lib.stat("quad_fallbacks") - f1 == 1
    ref = want.sum(dim=3, keepdim=True)
    assert float((base - ref).norm() / ref.norm()) <= 1e-5


def test_options_are_validated_and_sync_is_per_call():
    lib = capi.load()
    sc = scene(B=1)
    inp = sc.inputs(gr.to_cuda)
    bad = capi.make_call_options()
    bad.struct_size = 8
    with pytest.raises(capi.DvrError, match="struct_size"):
        render(lib, inp, 1, sc.H, sc.W, bad)
    bad = capi.make_call_options(volume_layout=5)
    with pytest.raises(capi.DvrError, match="volume_layout"):
        render(lib, inp, 1, sc.H, sc.W, bad)
    color = torch.empty(1, sc.H, sc.W, 4, device="cuda")
    term = torch.empty(1, sc.H, sc.W, dtype=torch.int32, device="cuda")
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):  # the default stream argument is torch's CURRENT stream
        lib.render_forward(inp, color, term, options=capi.make_call_options(sync=True))
        assert s.query()         # sync = 1: the call returned after its launches finished
    lib.trim_scratch()


def test_ctypes_path_rejects_wrong_dtype_and_device():
    lib = capi.load()
    sc = scene(B=1)
    inp = sc.inputs(gr.to_cuda)
    color = torch.empty(1, sc.H, sc.W, 4, device="cuda")
    term64 = torch.empty(1, sc.H, sc.W, dtype=torch.int64, device="cuda")
    with pytest.raises(TypeError, match="terminationIndex"):
        lib.render_forward(inp, color, term64)
    term = torch.empty(1, sc.H, sc.W, dtype=torch.int32, device="cuda")
    with pytest.raises(ValueError, match="color"):
        lib.render_forward(inp, color.cpu(), term)
    inp.volume = inp.volume.double()
    with pytest.raises(TypeError, match="volume"):
        lib.render_forward(inp, color, term)
    inp.volume = inp.volume.float().cpu()
    with pytest.raises(ValueError, match="no CPU path"):
        lib.render_forward(inp, color, term)
