"""Runs the CUDA product through the C++ host layer + pybind11 module ``pyrenderer`` (the
reference's Python API, bindings/bindings.cpp:382-511) on a tests.scenes.Scene. Same functions as
gpu_runner (which goes through the ctypes mirror of the C-ABI)."""
from __future__ import annotations

import numpy as np
import torch

from diffdvr_b200 import capi, load_pyrenderer
from gpu_runner import to_cuda

pyr = load_pyrenderer()

_FILTER = {capi.FILTER_NEAREST: pyr.VolumeFilterMode.Nearest, capi.FILTER_TRILINEAR: pyr.VolumeFilterMode.Trilinear,
           capi.FILTER_PRESHADED: pyr.VolumeFilterMode.Preshaded}
_CAMERA = {capi.CAMERA_RAY_START_DIR: pyr.CameraMode.RayStartDir,
           capi.CAMERA_INVERSE_VIEW_MATRIX: pyr.CameraMode.InverseViewMatrix,
           capi.CAMERA_REFERENCE_FRAME: pyr.CameraMode.ReferenceFrame}
_TF = {capi.TF_IDENTITY: pyr.TFMode.Identity, capi.TF_TEXTURE: pyr.TFMode.Texture, capi.TF_LINEAR: pyr.TFMode.Linear,
       capi.TF_GAUSSIAN: pyr.TFMode.Gaussian, capi.TF_GAUSSIAN_LOG: pyr.TFMode.GaussianLog,
       capi.TF_PRESHADED: pyr.TFMode.Preshaded}
_BLEND = {capi.BLEND_BEER_LAMBERT: pyr.BlendMode.BeerLambert, capi.BLEND_ALPHA: pyr.BlendMode.Alpha}


def make_inputs(scene, device="cuda"):
    """pyrenderer.RendererInputs for a Scene (tensors on ``device``)."""
    def t(a):
        return to_cuda(a) if device == "cuda" else torch.from_numpy(np.ascontiguousarray(a))

    def vec_or_tensor(a):
        return t(a) if isinstance(a, np.ndarray) else pyr.real3(*[float(v) for v in a])

    i = pyr.RendererInputs()
    i.screen_size = pyr.int2(scene.W, scene.H)
    i.volume_filter_mode = _FILTER[scene.filter_mode]
    i.volume = t(scene.volume)
    i.box_min = vec_or_tensor(scene.box_min)
    i.box_size = vec_or_tensor(scene.box_size)
    i.camera_mode = _CAMERA[scene.camera_mode]
    if scene.camera_mode == capi.CAMERA_RAY_START_DIR:
        i.camera = pyr.CameraPerPixelRays(t(scene.ray_start), t(scene.ray_dir))
    elif scene.camera_mode == capi.CAMERA_REFERENCE_FRAME:
        i.camera = pyr.CameraReferenceFrame(t(scene.camera_matrix), float(scene.fov_y))
    elif scene.camera_matrix.ndim == 3:
        i.camera = t(scene.camera_matrix)
    else:
        i.camera = pyr.mat4(np.ascontiguousarray(scene.camera_matrix, dtype=np.float32))
    i.step_size = t(scene.step_size) if isinstance(scene.step_size, np.ndarray) else float(scene.step_size)
    i.tf_mode = _TF[scene.tf_mode]
    i.tf = t(scene.tf)
    i.blend_mode = _BLEND[scene.blend_mode]
    i.blending_early_out = float(scene.early_out)
    return i


def forward(scene):
    B, H, W = scene.B, scene.H, scene.W
    out = pyr.RendererOutputs(torch.empty(B, H, W, 4, device="cuda"),
                              torch.empty(B, H, W, dtype=torch.int32, device="cuda"))
    pyr.Renderer.render_forward(make_inputs(scene), out)
    pyr.sync()
    return out.color.cpu().numpy(), out.terminationIndex.cpu().numpy()


def _adjoint_outputs(B, H, W, *, volume=None, step=False, camera=False, tf_shape=None, step_shape=None, delayed=False):
    a, out = pyr.AdjointOutputs(), {}
    if volume is not None:
        a.has_volume_derivatives = True
        a.adj_volume = out["adj_volume"] = torch.zeros_like(volume)
    if step:
        a.has_stepsize_derivatives = True
        a.adj_stepsize = out["adj_step_size"] = torch.zeros(step_shape or (B, H, W), device="cuda")
    if camera:
        a.has_camera_derivatives = True
        a.adj_camera_ray_start = out["adj_camera_ray_start"] = torch.zeros(B, H, W, 3, device="cuda")
        a.adj_camera_ray_dir = out["adj_camera_ray_dir"] = torch.zeros(B, H, W, 3, device="cuda")
    if tf_shape is not None:
        a.has_tf_derivatives = True
        a.tf_delayed_accumulation = delayed
        a.adj_tf = out["adj_tf"] = torch.zeros(tf_shape, device="cuda")
    return a, out


def adjoint(scene, color, term, adj_color, *, volume=False, step=False, camera=False, tf=False, step_shape=None,
            delayed=False):
    B, H, W = scene.B, scene.H, scene.W
    inp = make_inputs(scene)
    a, out = _adjoint_outputs(B, H, W, volume=inp.volume if volume else None, step=step, camera=camera,
                              tf_shape=scene.tf.shape if tf else None, step_shape=step_shape, delayed=delayed)
    fwd = pyr.RendererOutputs(torch.from_numpy(color).cuda(), torch.from_numpy(term).cuda())
    pyr.Renderer.render_adjoint(inp, fwd, torch.from_numpy(adj_color).cuda(), a)
    pyr.sync()
    return {k: v.cpu().numpy() for k, v in out.items()}


def _settings(D, d_stepsize, d_ray_start, d_ray_dir, d_tf):
    s = pyr.ForwardDifferencesSettings()
    s.D = int(D)
    s.d_stepsize = int(d_stepsize)
    s.d_rayStart = pyr.int3(*[int(v) for v in d_ray_start])
    s.d_rayDir = pyr.int3(*[int(v) for v in d_ray_dir])
    if d_tf is not None:
        s.d_tf = torch.from_numpy(d_tf).cuda()
        s.has_tf_derivatives = True
    return s


def forward_gradients(scene, D, *, d_stepsize=-1, d_ray_start=(-1, -1, -1), d_ray_dir=(-1, -1, -1), d_tf=None):
    B, H, W = scene.B, scene.H, scene.W
    out = pyr.RendererOutputs(torch.empty(B, H, W, 4, device="cuda"),
                              torch.empty(B, H, W, dtype=torch.int32, device="cuda"))
    grad = torch.zeros(B, H, W, D, 4, device="cuda")
    pyr.Renderer.render_forward_gradients(make_inputs(scene), _settings(D, d_stepsize, d_ray_start, d_ray_dir, d_tf),
                                          out, grad)
    pyr.sync()
    return out.color.cpu().numpy(), out.terminationIndex.cpu().numpy(), grad.cpu().numpy()


def forward_variables_to_gradients(fv, adj_color, D, *, tf_shape=None, d_stepsize=-1, d_ray_start=(-1, -1, -1),
                                   d_ray_dir=(-1, -1, -1), d_tf=None):
    B, H, W = fv.shape[:3]
    cam = any(v >= 0 for v in tuple(d_ray_start) + tuple(d_ray_dir))
    a, out = _adjoint_outputs(B, H, W, step=d_stepsize >= 0, camera=cam, tf_shape=tf_shape if d_tf is not None else None)
    pyr.Renderer.forward_variables_to_gradients(torch.from_numpy(fv).cuda(), torch.from_numpy(adj_color).cuda(),
                                                _settings(D, d_stepsize, d_ray_start, d_ray_dir, d_tf), a)
    pyr.sync()
    return {k: v.cpu().numpy() for k, v in out.items()}
