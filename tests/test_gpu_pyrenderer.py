"""GPU: the reference's Python API (`pyrenderer`, C++ host layer -> C-ABI -> sm_100a kernels) against
the golden fixtures of the reference's compiled host kernels, against the ctypes path (same library,
so bit-identical), plus the error behaviour and stream semantics of the reference's host layer."""
import os

import numpy as np
import pytest
import torch

import golden_cases
import gpu_runner as gr
import pyr_runner as pr
import scenes
from diffdvr_b200 import capi

pytestmark = pytest.mark.gpu
pyr = pr.pyr
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = {name: (sc, flags, fg) for name, sc, flags, fg in golden_cases.cases()}


@pytest.mark.parametrize("name", sorted(CASES))
def test_pyrenderer_matches_golden_reference_outputs(name):
    sc, flags, fg = CASES[name]
    gold = np.load(os.path.join(GOLDEN, f"ref_{name}.npz"))
    n0 = pyr.launch_count()
    pyr.set_option("compat_stepsize_overwrite", 1)
    try:
        out = golden_cases.run(None, sc, flags, fg, runner=pr, forward_from=(gold["color"], gold["term"]))
        via_ctypes = golden_cases.run(None, sc, flags, fg, runner=gr, forward_from=(gold["color"], gold["term"]),
                                      runner_kwargs=dict(options=capi.make_call_options(compat_stepsize_overwrite=True)))
    finally:
        pyr.set_option("compat_stepsize_overwrite", 0)
    assert pyr.launch_count() > n0                      # the CUDA library did the work
    assert scenes.rel_l2(out["color"], gold["color"]) <= 1e-5
    assert np.array_equal(out["color"], via_ctypes["color"]) and np.array_equal(out["term"], via_ctypes["term"])
    for k in gold.files:
        if k in ("adj_volume", "adj_tf"):
            assert scenes.rel_l2(out[k], gold[k]) <= 1e-4, (k, scenes.rel_l2(out[k], gold[k]))
        elif k.startswith("adj_"):                      # per-ray outputs: no atomics, same kernels -> identical
            assert np.array_equal(out[k], via_ctypes[k]), k
    if "fg_gradients" in gold.files:
        assert np.array_equal(out["fg_gradients"], via_ctypes["fg_gradients"])


def test_forward_variables_to_gradients_through_pyrenderer():
    sc = scenes.make_scene(n=16, W=12, H=10, B=2, camera_mode=capi.CAMERA_RAY_START_DIR, tf_mode=capi.TF_TEXTURE,
                           R=4, seed=5, opacity=4.0)
    kw = dict(d_stepsize=6, d_ray_start=(0, 1, 2), d_ray_dir=(3, 4, 5))
    _, _, fv = pr.forward_gradients(sc, 7, **kw)
    adj_color = np.random.RandomState(3).randn(2, 10, 12, 4).astype(np.float32)
    a = pr.forward_variables_to_gradients(fv, adj_color, 7, **kw)
    b = gr.forward_variables_to_gradients(fv, adj_color, 7, **kw)
    want = np.einsum("bhwc,bhwdc->bhwd", adj_color.astype(np.float64), fv.astype(np.float64))
    assert scenes.rel_l2(a["adj_step_size"], want[..., 6]) <= 1e-6
    assert scenes.rel_l2(a["adj_camera_ray_start"], want[..., 0:3]) <= 1e-6
    for k in a:
        assert np.array_equal(a[k], b[k]), k


def test_error_behaviour_of_the_host_layer():
    sc = scenes.make_scene(n=8, W=6, H=4, R=4)
    inp = pr.make_inputs(sc)
    good = pyr.RendererOutputs(torch.empty(1, 4, 6, 4, device="cuda"), torch.empty(1, 4, 6, dtype=torch.int32, device="cuda"))
    # wrong output shape: the reference's CHECK_SIZE wording (renderer/pytorch_utils.h:9)
    bad = pyr.RendererOutputs(torch.empty(1, 4, 5, 4, device="cuda"), good.terminationIndex)
    with pytest.raises(RuntimeError, match="color must have 6 entries at dimension 2, but has 5 entries"):
        pyr.Renderer.render_forward(inp, bad)
    # wrong dtypes
    with pytest.raises(RuntimeError, match="terminationIndex must be of type"):
        pyr.Renderer.render_forward(inp, pyr.RendererOutputs(good.color, good.color[..., 0]))
    dbl = inp.clone()
    dbl.tf = inp.tf.double()
    with pytest.raises(RuntimeError, match="tf must be of type"):
        pyr.Renderer.render_forward(dbl, good)
    # mixed devices
    mixed = inp.clone()
    mixed.tf = inp.tf.cpu()
    with pytest.raises(RuntimeError, match="All tensors must be on the same device, but tf"):
        pyr.Renderer.render_forward(mixed, good)
    # camera variant does not match the mode (renderer.cpp:127,152,182)
    cam = inp.clone()
    cam.camera_mode = pyr.CameraMode.RayStartDir
    with pytest.raises(RuntimeError, match="does not hold the first variant"):
        pyr.Renderer.render_forward(cam, good)
    # no adjoint variable requested (renderer.cpp:711-716)
    pyr.Renderer.render_forward(inp, good)
    with pytest.raises(RuntimeError, match="At least one adjoint variable must be activated"):
        pyr.Renderer.render_adjoint(inp, good, torch.zeros_like(good.color), pyr.AdjointOutputs())
    # forward variables cannot carry volume derivatives (renderer.cpp:541-542)
    a = pyr.AdjointOutputs()
    a.has_volume_derivatives = True
    a.adj_volume = torch.zeros_like(inp.volume)
    with pytest.raises(RuntimeError, match="volume derivatives are not supported"):
        pyr.Renderer.forward_variables_to_gradients(torch.zeros(1, 4, 6, 1, 4, device="cuda"),
                                                    torch.zeros(1, 4, 6, 4, device="cuda"),
                                                    pyr.ForwardDifferencesSettings(), a)
    # the library is still usable after the failures
    pyr.Renderer.render_forward(inp, good)
    pyr.sync()
    assert torch.isfinite(good.color).all()


def test_launches_go_to_the_current_torch_stream_and_timer_works():
    sc = scenes.make_scene(n=24, W=64, H=48, B=2, R=8, seed=2)
    ref_color, ref_term = pr.forward(sc)
    inp = pr.make_inputs(sc)
    side = torch.cuda.Stream()
    out = pyr.RendererOutputs(torch.zeros(2, 48, 64, 4, device="cuda"), torch.zeros(2, 48, 64, dtype=torch.int32, device="cuda"))
    torch.cuda.synchronize()
    with torch.cuda.stream(side):
        pyr.Renderer.render_forward(inp, out)
        done = torch.cuda.Event()
        done.record(side)          # ordered after the launch only if the launch went to `side`
    done.synchronize()
    assert np.array_equal(out.color.cpu().numpy(), ref_color) and np.array_equal(out.terminationIndex.cpu().numpy(), ref_term)
    t = pyr.GpuTimer()
    t.start()
    pyr.Renderer.render_forward(inp, out)
    t.stop()
    assert 0.0 < t.elapsed_ms() < 1000.0


# ---- Camera / compareToImage through the host layer --------------------------
def test_generate_rays_autograd_matches_capi_and_oracle():
    import camera_cases as cc
    import oracle_loader as ol
    case = cc.ray_cases()["rays_packed"]
    want = cc.run_rays(ol.port(32), case)
    vp = torch.from_numpy(want["viewport"]).cuda().requires_grad_(True)
    start, direction = pyr.Camera.generate_rays(vp, case["fovY"], case["W"], case["H"])
    assert start.shape == (case["B"], case["H"], case["W"], 3)
    assert np.array_equal(start.detach().cpu().numpy(), want["ray_start"])
    assert scenes.rel_l2(direction.detach().cpu().numpy(), want["ray_dir"]) <= 1e-6
    gs, gd = torch.from_numpy(want["adj_ray_start"]).cuda(), torch.from_numpy(want["adj_ray_dir"]).cuda()
    ((start * gs).sum() + (direction * gd).sum()).backward()
    assert scenes.rel_l2(vp.grad.cpu().numpy(), want["adj_viewport"]) <= 1e-4
    # only one of the two outputs reaches the loss
    vp2 = vp.detach().clone().requires_grad_(True)
    _, d2 = pyr.Camera.generate_rays(vp2, case["fovY"], case["W"], case["H"])
    (d2 * gd).sum().backward()
    assert torch.isfinite(vp2.grad).all() and (vp2.grad[:, 0] == 0).all()  # eye does not influence directions


def test_sphere_camera_to_adjoint_chain():
    """run_entropy2d-style chain: (yaw, pitch) -> viewport_from_sphere -> generate_rays -> render -> loss.
    d loss / d (yaw, pitch) = <adj_rays from render_adjoint(camera), d rays / d (yaw, pitch)>; the second factor
    is checked by central differences of the (smooth) ray generator, the first is pinned by the adjoint tests.
    (Differencing the RENDER itself is no oracle here: like the reference, the adjoint treats the box entry
    parameter tmin as a constant, renderer_kernels.cuh:336-345.)"""
    sc = scenes.make_scene(n=24, W=32, H=24, B=1, camera_mode=capi.CAMERA_RAY_START_DIR, tf_mode=capi.TF_GAUSSIAN, R=3,
                           seed=9, opacity=6.0)
    fov, W, H = 0.7, sc.W, sc.H
    center = torch.zeros(1, 3, device="cuda")
    dist = torch.tensor([[1.6]], device="cuda")
    weight = torch.rand(1, H, W, 4, generator=torch.Generator().manual_seed(0)).cuda()

    def rays(yaw, pitch):
        vp = pyr.Camera.viewport_from_sphere(center, yaw, pitch, dist, pyr.Orientation.Ym)
        return pyr.Camera.generate_rays(vp, fov, W, H)

    yaw = torch.tensor([[0.4]], device="cuda", requires_grad=True)
    pitch = torch.tensor([[0.3]], device="cuda", requires_grad=True)
    start, direction = rays(yaw, pitch)
    inp = pr.make_inputs(sc)
    inp.camera = pyr.CameraPerPixelRays(start.detach(), direction.detach())
    out = pyr.RendererOutputs(torch.empty(1, H, W, 4, device="cuda"),
                              torch.empty(1, H, W, dtype=torch.int32, device="cuda"))
    pyr.Renderer.render_forward(inp, out)
    assert out.color[..., 3].max().item() > 0.1  # the volume is in view
    a, o = pr._adjoint_outputs(1, H, W, camera=True)
    pyr.Renderer.render_adjoint(inp, out, weight, a)
    adj_s, adj_d = o["adj_camera_ray_start"], o["adj_camera_ray_dir"]
    assert adj_d.abs().max().item() > 0
    torch.autograd.backward([start, direction], [adj_s, adj_d])

    eps = 1e-3
    want = []
    for which in (0, 1):
        dy = eps if which == 0 else 0.0
        dp = eps if which == 1 else 0.0
        with torch.no_grad():
            s1, d1 = rays(yaw + dy, pitch + dp)
            s0, d0 = rays(yaw - dy, pitch - dp)
        want.append((((s1 - s0).double() / (2 * eps)) * adj_s.double()).sum().item() +
                    (((d1 - d0).double() / (2 * eps)) * adj_d.double()).sum().item())
    got = np.array([yaw.grad.item(), pitch.grad.item()])
    want = np.array(want)
    assert np.linalg.norm(got - want) <= 2e-3 * np.linalg.norm(want), (got, want)


def test_compare_to_image_through_host_layer():
    import camera_cases as cc
    import oracle_loader as ol
    case = cc.compare_cases()["compare_d6"]
    want = cc.run_compare(ol.port(32), case)
    rng = np.random.RandomState(cc.SEED + 1)
    B, W, H, D = case["B"], case["W"], case["H"], case["D"]
    color = torch.from_numpy(rng.uniform(0, 1, (B, H, W, 4)).astype(np.float32)).cuda()
    reference = torch.from_numpy(rng.uniform(0, 1, (B, H, W, 4)).astype(np.float32)).cuda()
    grads = torch.from_numpy(rng.standard_normal((B, H, W, D, 4)).astype(np.float32)).cuda()
    diff, gout = torch.empty(B, H, W, 1, device="cuda"), torch.empty(B, H, W, D, device="cuda")
    pyr.Renderer.compare_to_image(color, grads, reference, diff, gout, False)
    dr, gr_ = torch.empty(1, device="cuda"), torch.empty(D, device="cuda")
    pyr.Renderer.compare_to_image(color, grads, reference, dr, gr_, True)
    assert scenes.rel_l2(diff.cpu().numpy(), want["difference"]) <= 1e-6
    assert scenes.rel_l2(gout.cpu().numpy(), want["gradients"]) <= 1e-6
    assert scenes.rel_l2(dr.cpu().numpy(), want["difference_reduced"]) <= 1e-4
    assert scenes.rel_l2(gr_.cpu().numpy(), want["gradients_reduced"]) <= 1e-4
    with pytest.raises(RuntimeError):  # wrong output shape for the reduced variant
        pyr.Renderer.compare_to_image(color, grads, reference, diff, gout, True)


def test_preshade_volume_autograd_and_preshaded_render():
    """TFUtils.preshade_volume: values and density gradient against the oracle; the result feeds the pre-shaded
    marcher (train_volume.py --preshaded pipeline)."""
    import camera_cases as cc
    import oracle_loader as ol
    case = cc.preshade_cases()["preshade_gaussian"]
    want = cc.run_preshade(ol.port(32), case)
    density = torch.from_numpy(want["volume"]).cuda().requires_grad_(True)
    tf = torch.from_numpy(np.ascontiguousarray(want["tf"])).cuda()
    color = pyr.TFUtils.preshade_volume(density, tf, pyr.TFMode.Gaussian)
    assert color.shape == (4,) + density.shape[1:]
    assert scenes.rel_l2(color.detach().cpu().numpy(), want["color"]) <= 1e-5
    (color * torch.from_numpy(want["adj_color"]).cuda()).sum().backward()
    assert scenes.rel_l2(density.grad.cpu().numpy(), want["adj_volume"]) <= 1e-4
    # pre-shaded render of the colour volume
    i = pyr.RendererInputs()
    i.screen_size = pyr.int2(32, 24)
    i.volume_filter_mode = pyr.VolumeFilterMode.Preshaded
    i.volume = color.detach()
    i.box_min = pyr.real3(-0.5, -0.5, -0.5)
    i.box_size = pyr.real3(1, 1, 1)
    i.camera_mode = pyr.CameraMode.ReferenceFrame
    vp = pyr.Camera.viewport_from_sphere(torch.zeros(1, 3, device="cuda"), torch.tensor([[0.5]], device="cuda"),
                                         torch.tensor([[0.3]], device="cuda"), torch.tensor([[1.8]], device="cuda"),
                                         pyr.Orientation.Ym)
    i.camera = pyr.CameraReferenceFrame(vp, 0.6)
    i.step_size = 0.02
    i.tf_mode = pyr.TFMode.Preshaded
    i.tf = tf
    i.blend_mode = pyr.BlendMode.BeerLambert
    out = pyr.RendererOutputs(torch.empty(1, 24, 32, 4, device="cuda"),
                              torch.empty(1, 24, 32, dtype=torch.int32, device="cuda"))
    pyr.Renderer.render_forward(i, out)
    pyr.sync()
    assert torch.isfinite(out.color).all() and out.color[..., 3].max().item() > 0.05
