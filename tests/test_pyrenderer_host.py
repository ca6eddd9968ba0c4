"""CPU: the C++ host layer + pybind11 module `pyrenderer` load without a GPU, expose the renderer
surface of the reference's module (bindings/bindings.cpp:143-232, 306-511, 554-558) and reject what
the product does not do (CPU tensors: there is no CPU back-end). No compute calls here."""
import numpy as np
import pytest
import torch

import pyr_runner
import scenes
from diffdvr_b200 import capi, load_pyrenderer

pyr = load_pyrenderer()


def test_module_surface_matches_the_reference_bindings():
    # names as registered by the reference's PYBIND11_MODULE for the renderer path
    functions = ["use_double_precision", "reload_kernels", "set_cuda_cache_dir", "disable_cuda_cache",
                 "set_cuda_sync_mode", "set_cuda_debug_mode", "set_kernel_cache_file", "cleanup", "sync"]
    classes = {"real3": ["x", "y", "z"], "real4": ["x", "y", "z", "w"], "int2": ["x", "y"], "int3": ["x", "y", "z"],
               "CameraPerPixelRays": ["ray_start", "ray_dir"], "CameraReferenceFrame": ["viewport", "fov_y_radians"],
               "mat4": ["to_torch"],
               "RendererInputs": ["clone", "screen_size", "volume", "volume_filter_mode", "box_min", "box_size",
                                  "camera_mode", "camera", "step_size", "tf_mode", "tf", "blend_mode",
                                  "blending_early_out"],
               "RendererOutputs": ["color", "terminationIndex"],
               "ForwardDifferencesSettings": ["D", "d_stepsize", "d_rayStart", "d_rayDir", "d_tf", "has_tf_derivatives"],
               "AdjointOutputs": ["has_volume_derivatives", "adj_volume", "has_stepsize_derivatives", "adj_stepsize",
                                  "has_camera_derivatives", "adj_camera_ray_start", "adj_camera_ray_dir",
                                  "has_tf_derivatives", "tf_delayed_accumulation", "adj_tf"],
               "Renderer": ["render_forward", "render_forward_gradients", "render_adjoint",
                            "forward_variables_to_gradients"],
               "GpuTimer": ["start", "stop", "elapsed_ms"]}
    for f in functions:
        assert callable(getattr(pyr, f)), f
    for c, attrs in classes.items():
        for a in attrs:
            assert hasattr(getattr(pyr, c), a), (c, a)
    assert pyr.use_double_precision() == 0


def test_enum_names_and_numeric_values():
    # renderer/renderer_settings.cuh:81-146
    want = {"VolumeFilterMode": dict(Nearest=0, Trilinear=1, Preshaded=3),
            "CameraMode": dict(RayStartDir=0, InverseViewMatrix=1, ReferenceFrame=2),
            "TFMode": dict(Identity=0, Texture=1, Linear=2, Gaussian=3, Preshaded=5, GaussianLog=6),
            "BlendMode": dict(BeerLambert=0, Alpha=1)}
    for enum, values in want.items():
        for name, v in values.items():
            assert int(getattr(getattr(pyr, enum), name)) == v


def test_variant_fields_accept_scalars_and_tensors():
    i = pyr.RendererInputs()
    assert abs(i.blending_early_out - (1 - 1e-5)) < 1e-7          # renderer.h:63
    i.box_min = pyr.real3(1, 2, 3)
    assert (i.box_min.x, i.box_min.y, i.box_min.z) == (1, 2, 3)
    i.box_min = torch.zeros(2, 3)
    assert tuple(i.box_min.shape) == (2, 3)
    i.step_size = 0.25
    assert i.step_size == 0.25
    i.step_size = torch.ones(1, 1, 1)
    assert isinstance(i.step_size, torch.Tensor)
    i.camera = pyr.CameraReferenceFrame(torch.eye(3)[None], 0.5)
    assert i.camera.fov_y_radians == 0.5
    i.camera = pyr.CameraPerPixelRays(torch.zeros(1, 2, 2, 3), torch.ones(1, 2, 2, 3))
    assert float(i.camera.ray_dir.sum()) == 12
    j = i.clone()
    j.step_size = 1.0
    assert isinstance(i.step_size, torch.Tensor)
    s = pyr.ForwardDifferencesSettings()
    assert s.d_stepsize == -1 and (s.d_rayStart.x, s.d_rayDir.z) == (-1, -1) and not s.has_tf_derivatives


def test_mat4_buffer_round_trip():
    a = np.arange(16, dtype=np.float32).reshape(4, 4)
    m = pyr.mat4(a)
    assert np.array_equal(np.array(m, copy=False), a)
    assert torch.equal(m.to_torch(), torch.from_numpy(a)[None])
    with pytest.raises(RuntimeError):
        pyr.mat4(np.zeros((3, 4), np.float32))
    with pytest.raises(RuntimeError):
        pyr.mat4(np.zeros((4, 4), np.float64))


def test_cpu_tensors_are_rejected_there_is_no_cpu_backend():
    sc = scenes.make_scene(n=8, W=4, H=4, R=4)
    inp = pyr_runner.make_inputs(sc, device="cpu")
    out = pyr.RendererOutputs(torch.zeros(1, 4, 4, 4), torch.zeros(1, 4, 4, dtype=torch.int32))
    with pytest.raises(RuntimeError, match="no CPU back-end"):
        pyr.Renderer.render_forward(inp, out)
    a = pyr.AdjointOutputs()
    a.has_tf_derivatives = True
    a.adj_tf = torch.zeros(1, 4, 4)
    with pytest.raises(RuntimeError, match="no CPU back-end"):
        pyr.Renderer.render_adjoint(inp, out, torch.zeros(1, 4, 4, 4), a)
    with pytest.raises(RuntimeError, match="no CPU back-end"):
        pyr.Renderer.forward_variables_to_gradients(torch.zeros(1, 4, 4, 2, 4), torch.zeros(1, 4, 4, 4),
                                                    pyr.ForwardDifferencesSettings(), a)
    u = pyr.RendererInputs()
    with pytest.raises(RuntimeError, match="volume is not defined"):
        pyr.Renderer.render_forward(u, out)


def test_options_and_control_shims():
    # options of the static facade (passed to the C-ABI per call) ...
    assert pyr.get_option("compat_stepsize_overwrite") == 0 and pyr.get_option("volume_layout") == -1
    assert pyr.get_option("volume_cache") == 1
    pyr.set_option("compat_stepsize_overwrite", 1)
    assert pyr.get_option("compat_stepsize_overwrite") == 1
    pyr.set_option("compat_stepsize_overwrite", 0)
    # ... and the library's one process-wide switch: one library behind both front ends
    pyr.set_option("sync_mode", 1)
    assert capi.load().get_option("sync_mode") == 1
    pyr.set_option("sync_mode", 0)
    assert pyr.stat("launches") == pyr.launch_count()
    with pytest.raises(RuntimeError):
        pyr.set_option("no_such_option", 1)
    pyr.set_cuda_sync_mode(True)
    assert pyr.get_cuda_sync_mode()
    pyr.set_cuda_sync_mode(False)
    for f in (pyr.reload_kernels, pyr.disable_cuda_cache, pyr.cleanup):
        f()
    pyr.set_cuda_cache_dir("/tmp/x")
    pyr.set_kernel_cache_file("k")
    pyr.set_cuda_debug_mode(False)
    assert pyr.has_variant(pyr.VolumeFilterMode.Trilinear, pyr.CameraMode.ReferenceFrame, pyr.TFMode.Texture,
                           pyr.BlendMode.BeerLambert)


# ---- Camera (bindings/bindings.cpp:514-551) ----------------------------------------------------------
def test_camera_surface():
    for name in ("Xp", "Xm", "Yp", "Ym", "Zp", "Zm"):
        assert hasattr(pyr.Orientation, name)
    for f in ("compute_matrix", "generate_rays", "viewport_from_lookat", "viewport_from_sphere"):
        assert callable(getattr(pyr.Camera, f)), f
    assert callable(pyr.Renderer.compare_to_image)


def test_viewport_from_lookat_matches_numpy_restatement():
    import camera_cases as cc
    rng = np.random.RandomState(2)
    origin, look_at = rng.uniform(-2, 2, (4, 3)).astype(np.float32), rng.uniform(-0.2, 0.2, (4, 3)).astype(np.float32)
    up = np.tile(np.array([[0, 1, 0]], np.float32), (4, 1))
    got = pyr.Camera.viewport_from_lookat(torch.from_numpy(origin), torch.from_numpy(look_at), torch.from_numpy(up))
    want = cc.viewport_from_lookat(origin.astype(np.float64), look_at.astype(np.float64), up.astype(np.float64))
    assert got.shape == (4, 3, 3)
    assert np.allclose(got.numpy(), want, atol=1e-6)
    # right-handed-looking frame of the reference: right = normalize(front x up), up2 = normalize(right x front)
    front = look_at - origin
    front /= np.linalg.norm(front, axis=1, keepdims=True)
    assert np.allclose(np.cross(got[:, 1].numpy(), front), got[:, 2].numpy(), atol=1e-5)


def test_viewport_from_sphere_semantics_and_autograd():
    # pytests/diffdvr/utils.py + renderer/camera.cpp:141-180: Ym orientation, yaw inverted, permutation (-1,-2,-3)
    center = torch.zeros(2, 3)
    yaw = torch.tensor([[0.3], [1.1]], requires_grad=True)
    pitch = torch.tensor([[0.2], [-0.4]], requires_grad=True)
    dist = torch.tensor([[1.5], [2.0]])
    vp = pyr.Camera.viewport_from_sphere(center, yaw, pitch, dist, pyr.Orientation.Ym)
    assert vp.shape == (2, 3, 3)
    # OrientationInvertPitch is false for every orientation: the pitch is always negated (camera.cpp:160)
    y, p, d = yaw.detach().numpy(), -pitch.detach().numpy(), dist.numpy()
    pos = np.concatenate([np.cos(p) * np.cos(y) * d, np.sin(p) * d, np.cos(p) * np.sin(y) * d], axis=1)
    assert np.allclose(vp[:, 0].detach().numpy(), -pos, atol=1e-6)  # origin = -pos for Ym (yaw * +1)
    assert np.allclose(np.linalg.norm(vp[:, 1].detach().numpy(), axis=1), 1, atol=1e-6)
    assert np.allclose((vp[:, 1] * vp[:, 2]).sum(1).detach().numpy(), 0, atol=1e-6)
    vp.square().sum().backward()
    assert yaw.grad is not None and pitch.grad is not None and torch.isfinite(yaw.grad).all()
    # Yp: yaw sign flipped (OrientationInvertYaw false -> * -1), identity permutation
    vp2 = pyr.Camera.viewport_from_sphere(center, yaw.detach(), pitch.detach(), dist, pyr.Orientation.Yp)
    pos2 = np.concatenate([np.cos(p) * np.cos(-y) * d, np.sin(p) * d, np.cos(p) * np.sin(-y) * d], axis=1)
    assert np.allclose(vp2[:, 0].numpy(), pos2, atol=1e-6)


def test_compute_matrix_is_the_inverse_view_projection():
    o, la, up = (-1.0, -0.5, 0.2), (0.0, 0.0, 0.0), (0.0, 0.0, 1.0)  # unittests/testConsistencyTF.cpp scene camera
    m = np.array(pyr.Camera.compute_matrix(pyr.real3(*o), pyr.real3(*la), pyr.real3(*up), 45.0, 128, 64, 0.1, 10.0))
    assert m.shape == (4, 4)
    o, la, up = (np.array(v, np.float64) for v in (o, la, up))
    f = (la - o) / np.linalg.norm(la - o)
    s = np.cross(up, f)
    s /= np.linalg.norm(s)
    u = np.cross(f, s)
    view = np.eye(4)
    view[0, :3], view[1, :3], view[2, :3] = s, u, f
    view[:3, 3] = [-s @ o, -u @ o, -f @ o]
    h = 1 / np.tan(np.radians(45.0) / 2)
    proj = np.zeros((4, 4))
    proj[0, 0], proj[1, 1] = h * 64 / 128, h
    proj[2, 2], proj[2, 3], proj[3, 2] = 10 / 9.9, -(10 * 0.1) / 9.9, 1
    assert np.allclose(m, np.linalg.inv(proj @ view), rtol=1e-5, atol=1e-5)
    # NDC (0, 0, z=0) unprojects onto the near plane straight ahead of the eye (matmul_proj semantics)
    w = m @ np.array([0, 0, 0, 1.0])
    assert np.allclose(w[:3] / w[3], o + 0.1 * f, atol=1e-5)


def test_generate_rays_rejects_cpu_tensors():
    with pytest.raises(RuntimeError, match="no CPU back-end"):
        pyr.Camera.generate_rays(torch.zeros(1, 3, 3), 0.5, 8, 8)
    with pytest.raises(RuntimeError, match="no CPU back-end"):
        pyr.Renderer.compare_to_image(torch.zeros(1, 2, 2, 4), torch.zeros(1, 2, 2, 1, 4), torch.zeros(1, 2, 2, 4),
                                      torch.zeros(1, 2, 2, 1), torch.zeros(1, 2, 2, 1), False)


# ---- TFUtils (bindings/bindings.cpp:560-571) --------------------------------------------------------------
def test_lab_rgb_round_trip_and_known_values():
    # D65 white: Lab (100, 0, 0) <-> rgb (1, 1, 1); mid grey rgb 0.5 -> L* = 53.39
    w = pyr.lab_to_rgb(pyr.real3(100.0, 0.0, 0.0))
    assert np.allclose([w.x, w.y, w.z], 1.0, atol=2e-3)
    g = pyr.rgb_to_lab(pyr.real3(0.5, 0.5, 0.5))
    assert abs(g.x - 53.389) < 0.02 and abs(g.y) < 0.05 and abs(g.z) < 0.05
    rng = np.random.RandomState(0)
    for rgb in rng.uniform(0.05, 0.95, (20, 3)):
        lab = pyr.rgb_to_lab(pyr.real3(*[float(v) for v in rgb]))
        back = pyr.lab_to_rgb(lab)
        assert np.allclose([back.x, back.y, back.z], rgb, atol=2e-4)


def _assemble(colors_lab, cpos, opac, opos, lo=0.0, hi=1.0, scale=10.0, purge=False):
    pts = pyr.TFUtils.assemble_from_settings([pyr.real3(*[float(v) for v in c]) for c in colors_lab],
                                             [float(v) for v in cpos], [float(v) for v in opac],
                                             [float(v) for v in opos], lo, hi, scale, purge)
    return pts


def test_assemble_from_settings_interpolates_both_lists():
    colors = np.array([[30.0, 20.0, -40.0], [70.0, -30.0, 50.0], [90.0, 5.0, 5.0]])
    cpos, opac, opos = [0.1, 0.5, 0.9], [0.0, 0.8, 0.2, 0.6], [0.05, 0.3, 0.6, 0.95]
    pts = _assemble(colors, cpos, opac, opos, scale=10.0)
    pos = np.array([p.pos for p in pts])
    assert np.all(np.diff(pos) >= 0) and pos[0] == -1.0           # sorted, guard point at -1
    assert set(np.round(pos, 6)) >= set(np.round(cpos + opos, 6))  # every control point of either list
    cg, og = [-1.0] + cpos + [2.0], [-1.0] + opos + [2.0]
    lab_g = np.vstack([colors[:1], colors, colors[-1:]])
    op_g = [opac[0]] + opac + [opac[-1]]
    for p in pts:
        want_a = np.clip(np.interp(p.pos, og, op_g), 0, 1 - 1.2e-7) * 10.0
        assert abs(p.val.w - want_a) < 1e-5
        lab = [np.interp(p.pos, cg, lab_g[:, k]) for k in range(3)]
        rgb = pyr.lab_to_rgb(pyr.real3(*[float(v) for v in lab]))
        want = np.clip([rgb.x, rgb.y, rgb.z], 0, 1 - 1.2e-7)
        assert np.allclose([p.val.x, p.val.y, p.val.z], want, atol=1e-5)


def test_assemble_maps_density_range_and_purges_zero_opacity():
    colors = np.array([[50.0, 0.0, 0.0], [60.0, 10.0, 10.0]])
    pts = _assemble(colors, [0.0, 1.0], [1.0, 1.0], [0.0, 1.0], lo=0.2, hi=0.6, scale=1.0)
    pos = sorted(p.pos for p in pts)
    assert any(abs(v - 0.2) < 1e-6 for v in pos) and any(abs(v - 0.6) < 1e-6 for v in pos)
    # a long zero-opacity stretch with several colour points inside: purge drops the interior ones
    cpos = [0.1, 0.2, 0.3, 0.4, 0.9]
    cols = np.tile(np.array([[50.0, 0.0, 0.0]]), (5, 1))
    keep = _assemble(cols, cpos, [0.0, 0.0, 1.0], [0.0, 0.5, 1.0], purge=False)
    purged = _assemble(cols, cpos, [0.0, 0.0, 1.0], [0.0, 0.5, 1.0], purge=True)
    assert len(purged) < len(keep)


def test_piecewise_and_texture_tensors():
    pts = [pyr.TFPoint(-1.0, pyr.real4(0.1, 0.2, 0.3, 0.0)), pyr.TFPoint(0.25, pyr.real4(0.1, 0.2, 0.3, 0.0)),
           pyr.TFPoint(0.75, pyr.real4(0.9, 0.6, 0.3, 8.0)), pyr.TFPoint(2.0, pyr.real4(0.9, 0.6, 0.3, 8.0))]
    lin = pyr.TFUtils.get_piecewise_tensor(pts)
    assert lin.shape == (1, 4, 5) and lin.dtype == torch.float32 and not lin.is_cuda
    assert np.allclose(lin[0, :, 4].numpy(), [-1, 0.25, 0.75, 2]) and np.allclose(lin[0, 2, :4].numpy(), [0.9, 0.6, 0.3, 8])
    tex = pyr.TFUtils.get_texture_tensor(pts, 16)
    assert tex.shape == (1, 16, 4)
    dens = (np.arange(16) + 0.5) / 16
    want_a = np.interp(dens, [-1, 0.25, 0.75, 2], [0, 0, 8, 8])
    assert np.allclose(tex[0, :, 3].numpy(), want_a, atol=1e-5)
    assert np.allclose(tex[0, :, 0].numpy(), np.interp(dens, [-1, 0.25, 0.75, 2], [0.1, 0.1, 0.9, 0.9]), atol=1e-6)
    with pytest.raises(RuntimeError, match="no CPU back-end"):
        pyr.TFUtils.preshade_volume(torch.zeros(1, 2, 2, 2), tex, pyr.TFMode.Texture)
