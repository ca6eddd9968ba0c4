"""GPU: dvr_render_forward_noalloc (include/dvr_b200.h) -- the forward pass whose QUAD copy lives in a static device array of
the library, for callers that cannot allocate. Checked through the C-ABI against the existing entry point
(dvr_render_forward_ex, volume_layout = 1), on two calls in a row with DIFFERENT volumes of different sizes, so that the
second call would show anything the shared array kept from the first. Collected last
(file name)."""
import numpy as np
import pytest
import torch

import gpu_runner as gr
import scenes
from diffdvr_b200 import capi

pytestmark = pytest.mark.gpu

IMAGE_TOL = 1e-5  # rel-L2, the image tolerance of tests/test_gpu_parity.py (the two paths run the same kernels)


def outputs(sc):
    color = torch.full((sc.B, sc.H, sc.W, 4), float("nan"), device="cuda")
    term = torch.full((sc.B, sc.H, sc.W), -7, dtype=torch.int32, device="cuda")
    return color, term


def compare(lib, sc, volume_per_view=False):
    inp = sc.inputs(gr.to_cuda)
    if volume_per_view:   # B volumes: B * (X + 1) * Y * Z entries of the array
        inp.volume = torch.cat([inp.volume * (1.0 - 0.3 * b) if b % 2 == 0 else inp.volume.flip(1) * 0.8
                                for b in range(sc.B)]).contiguous()
    want_c, want_t = outputs(sc)
    lib.render_forward(inp, want_c, want_t, options=capi.make_call_options(volume_layout=1))
    got_c, got_t = outputs(sc)
    lib.render_forward_noalloc(inp, got_c, got_t)
    torch.cuda.synchronize()
    assert torch.isfinite(got_c).all() and float(want_c.abs().max()) > 0.0
    err = scenes.rel_l2(got_c.cpu().numpy(), want_c.cpu().numpy())
    assert err <= IMAGE_TOL, err
    assert torch.equal(got_t, want_t)
    return got_c


@pytest.mark.parametrize("tf_mode", [capi.TF_TEXTURE, capi.TF_GAUSSIAN])
def test_two_calls_in_a_row_match_the_allocating_entry_point(tf_mode):
    lib = capi.load()
    # a batch of two volumes first (fills more of the array), then ONE smaller volume with other contents
    first = scenes.make_scene(n=40, W=56, H=44, B=2, tf_mode=tf_mode, R=8, seed=11, opacity=3.0)
    second = scenes.make_scene(n=24, W=37, H=29, B=1, tf_mode=tf_mode, R=8, seed=12, opacity=2.0,
                               camera_mode=capi.CAMERA_INVERSE_VIEW_MATRIX)
    a = compare(lib, first, volume_per_view=True)
    b = compare(lib, second)
    # and the first again: the array now holds the second call's copy
    a2 = compare(lib, first, volume_per_view=True)
    assert torch.equal(a, a2) and a.shape != b.shape


def test_largest_volume_is_accepted_and_one_slice_more_is_rejected():
    lib = capi.load()
    sc = scenes.make_scene(n=16, W=32, H=24, B=1, tf_mode=capi.TF_TEXTURE, R=8, seed=5, opacity=3.0)
    inp = sc.inputs(gr.to_cuda)
    color, term = outputs(sc)
    g = torch.Generator(device="cuda").manual_seed(1)
    inp.volume = torch.rand(1, 128, 128, 128, device="cuda", generator=g)          # (128 + 1) * 128 * 128: the limit
    assert (128 + 1) * 128 * 128 == capi.NOALLOC_MAX_QUADS
    want_c, want_t = outputs(sc)
    lib.render_forward(inp, want_c, want_t, options=capi.make_call_options(volume_layout=1))
    lib.render_forward_noalloc(inp, color, term)
    torch.cuda.synchronize()
    assert scenes.rel_l2(color.cpu().numpy(), want_c.cpu().numpy()) <= IMAGE_TOL and torch.equal(term, want_t)
    n0 = lib.launch_count()
    inp.volume = torch.rand(1, 129, 128, 128, device="cuda", generator=g)
    with pytest.raises(Exception, match="DVR_NOALLOC_MAX_QUADS"):
        lib.render_forward_noalloc(inp, color, term)
    assert lib.launch_count() == n0   # rejected before any launch


def test_filters_without_scratch_run_unchanged():
    lib = capi.load()
    sc = scenes.make_scene(n=20, W=33, H=21, B=1, tf_mode=capi.TF_LINEAR, R=6, seed=7, opacity=3.0,
                           filter_mode=capi.FILTER_NEAREST)
    inp = sc.inputs(gr.to_cuda)
    want_c, want_t = outputs(sc)
    lib.render_forward(inp, want_c, want_t)
    got_c, got_t = outputs(sc)
    b0 = lib.stat("quad_builds")
    lib.render_forward_noalloc(inp, got_c, got_t)
    torch.cuda.synchronize()
    assert lib.stat("quad_builds") == b0
    assert torch.equal(got_c, want_c) and torch.equal(got_t, want_t)
