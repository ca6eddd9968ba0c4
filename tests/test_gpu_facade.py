"""GPU: renderer::RendererCuda over the C-ABI (diffdvr_b200/libdvr_renderer_cuda.so), driven the way the reference's
host layer drives its back-end (tests/csrc/facade_shim.cpp fills the reference's packed accessors). Same library
underneath, so every result must be bit-identical to the direct C-ABI call -- atomically accumulated gradients
(adj_volume with a non-QUAD layout, adj_tf) within the gradient tolerance."""
import ctypes

import numpy as np
import pytest
import torch

import facade_loader as fl
import gpu_runner as gr
import scenes
from diffdvr_b200 import capi

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def shim():
    return fl.load()


def _scene(camera_mode, tf_mode, **kw):
    return scenes.make_scene(n=20, W=24, H=18, B=2, camera_mode=camera_mode, tf_mode=tf_mode, R=6, seed=4, opacity=5.0,
                             tensor_scalars=True, **kw)


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


@pytest.mark.parametrize("camera_mode,tf_mode", [(capi.CAMERA_REFERENCE_FRAME, capi.TF_TEXTURE),
                                                 (capi.CAMERA_RAY_START_DIR, capi.TF_LINEAR),
                                                 (capi.CAMERA_INVERSE_VIEW_MATRIX, capi.TF_GAUSSIAN)])
def test_forward_and_adjoint_are_the_c_abi_results(shim, camera_mode, tf_mode):
    lib = capi.load()
    sc = _scene(camera_mode, tf_mode)
    B, H, W = sc.B, sc.H, sc.W
    want_c, want_t = gr.forward(sc, lib)

    inp = sc.inputs(gr.to_cuda)
    c_in = inp.to_c()
    color = torch.empty(B, H, W, 4, device="cuda")
    term = torch.empty(B, H, W, dtype=torch.int32, device="cuda")
    out = capi.make_outputs(color, term)
    n0 = lib.launch_count()
    assert shim.shim_render_forward(ctypes.byref(c_in), ctypes.byref(out), _stream()) == 1, shim.shim_last_error()
    torch.cuda.synchronize()
    assert lib.launch_count() > n0
    assert np.array_equal(color.cpu().numpy(), want_c) and np.array_equal(term.cpu().numpy(), want_t)

    adj_color = np.random.RandomState(1).randn(B, H, W, 4).astype(np.float32)
    flags = dict(volume=True, tf=True, step=True, camera=camera_mode == capi.CAMERA_RAY_START_DIR)
    want = gr.adjoint(sc, want_c, want_t, adj_color, lib=lib, options=capi.make_call_options(compat_stepsize_overwrite=True),
                      **flags)
    bufs = dict(adj_volume=torch.zeros_like(inp.volume), adj_tf=torch.zeros(sc.tf.shape, device="cuda"),
                adj_step_size=torch.zeros(B, H, W, device="cuda"))
    if flags["camera"]:
        bufs["adj_camera_ray_start"] = torch.zeros(B, H, W, 3, device="cuda")
        bufs["adj_camera_ray_dir"] = torch.zeros(B, H, W, 3, device="cuda")
    adj = capi.make_adjoint_outputs(**bufs)
    # DvrTensor views do not own their tensors: keep the device copies alive until the call has run
    color_in, term_in, adj_color_in = torch.from_numpy(want_c).cuda(), torch.from_numpy(want_t).cuda(), \
        torch.from_numpy(adj_color).cuda()
    fwd = capi.make_outputs(color_in, term_in)
    ac = capi.tensor_view(adj_color_in)
    keep = (fwd, ac, color_in, term_in, adj_color_in)
    assert shim.shim_render_adjoint(ctypes.byref(c_in), ctypes.byref(fwd), ctypes.byref(ac), ctypes.byref(adj),
                                    _stream()) == 1, shim.shim_last_error()
    torch.cuda.synchronize()
    del keep
    for k, v in bufs.items():
        got = v.cpu().numpy()
        if k in ("adj_volume", "adj_tf"):
            assert scenes.rel_l2(got, want[k]) <= 1e-5, k
        else:
            assert np.array_equal(got, want[k]), k


def test_forward_mode_and_its_reduction(shim):
    lib = capi.load()
    sc = _scene(capi.CAMERA_RAY_START_DIR, capi.TF_TEXTURE)
    B, H, W = sc.B, sc.H, sc.W
    R, C = sc.tf.shape[1:]
    d_tf = (7 + np.arange(R * C, dtype=np.int32)).reshape(1, R, C)
    D = 7 + R * C
    kw = dict(d_stepsize=6, d_ray_start=(0, 1, 2), d_ray_dir=(3, 4, 5), d_tf=d_tf)
    want_c, want_t, want_g = gr.forward_gradients(sc, D, lib=lib, **kw)

    inp = sc.inputs(gr.to_cuda)
    c_in = inp.to_c()
    dtf = torch.from_numpy(d_tf).cuda()
    st = capi.make_fd_settings(D, 6, (0, 1, 2), (3, 4, 5), dtf, True)
    color = torch.empty(B, H, W, 4, device="cuda")
    term = torch.empty(B, H, W, dtype=torch.int32, device="cuda")
    grad = torch.zeros(B, H, W, D, 4, device="cuda")
    out, g = capi.make_outputs(color, term), capi.tensor_view(grad)
    assert shim.shim_render_forward_gradients(ctypes.byref(c_in), ctypes.byref(st), ctypes.byref(out), ctypes.byref(g),
                                              _stream()) == 1, shim.shim_last_error()
    torch.cuda.synchronize()
    assert np.array_equal(color.cpu().numpy(), want_c) and np.array_equal(term.cpu().numpy(), want_t)
    assert np.array_equal(grad.cpu().numpy(), want_g)

    adj_color = np.random.RandomState(2).randn(B, H, W, 4).astype(np.float32)
    want = gr.forward_variables_to_gradients(want_g, adj_color, D, tf_shape=sc.tf.shape, lib=lib, **kw)
    bufs = dict(adj_step_size=torch.zeros(B, H, W, device="cuda"), adj_camera_ray_start=torch.zeros(B, H, W, 3, device="cuda"),
                adj_camera_ray_dir=torch.zeros(B, H, W, 3, device="cuda"), adj_tf=torch.zeros(sc.tf.shape, device="cuda"))
    adj = capi.make_adjoint_outputs(**bufs)
    ac_t = torch.from_numpy(adj_color).cuda()
    ac = capi.tensor_view(ac_t)
    assert shim.shim_forward_variables_to_gradients(ctypes.byref(g), ctypes.byref(ac), ctypes.byref(st), ctypes.byref(adj),
                                                    _stream()) == 1, shim.shim_last_error()
    torch.cuda.synchronize()
    for k, v in bufs.items():
        if k == "adj_tf":
            assert scenes.rel_l2(v.cpu().numpy(), want[k]) <= 1e-5
        else:
            assert np.array_equal(v.cpu().numpy(), want[k]), k


def test_a_failed_call_returns_false_with_the_reason(shim):
    sc = _scene(capi.CAMERA_REFERENCE_FRAME, capi.TF_TEXTURE)
    inp = sc.inputs(gr.to_cuda)
    c_in = inp.to_c()
    color = torch.empty(2, 18, 23, 4, device="cuda")  # W is 24
    term = torch.empty(2, 18, 24, dtype=torch.int32, device="cuda")
    out = capi.make_outputs(color, term)
    assert shim.shim_render_forward(ctypes.byref(c_in), ctypes.byref(out), _stream()) == 0
    assert b"color" in shim.shim_last_error()
