"""Runs the CUDA product (through the C-ABI, ctypes) on a tests.scenes.Scene."""
from __future__ import annotations

import numpy as np
import torch

from diffdvr_b200 import capi


def to_cuda(a: np.ndarray):
    t = torch.from_numpy(np.ascontiguousarray(a)).cuda()
    if a.strides != np.ascontiguousarray(a).strides and a.ndim == 4:
        # preserve a non-contiguous (x-fastest) volume layout on the device
        order = np.argsort([-s for s in a.strides])
        t = torch.from_numpy(np.ascontiguousarray(a.transpose(order))).cuda().permute(*np.argsort(order))
    return t


def forward(scene, lib=None, options=None):
    lib = lib or capi.load()
    B, H, W = scene.B, scene.H, scene.W
    inp = scene.inputs(to_cuda)
    color = torch.empty(B, H, W, 4, device="cuda")
    term = torch.empty(B, H, W, dtype=torch.int32, device="cuda")
    lib.render_forward(inp, color, term, torch.cuda.current_stream().cuda_stream, options)
    torch.cuda.synchronize()
    return color.cpu().numpy(), term.cpu().numpy()


def adjoint(scene, color, term, adj_color, *, volume=False, step=False, camera=False, tf=False,
            step_shape=None, delayed=False, lib=None, options=None):
    lib = lib or capi.load()
    B, H, W = scene.B, scene.H, scene.W
    inp = scene.inputs(to_cuda)
    out = {}
    if volume:
        out["adj_volume"] = torch.zeros_like(inp.volume)
    if step:
        out["adj_step_size"] = torch.zeros(step_shape or (B, H, W), device="cuda")
    if camera:
        out["adj_camera_ray_start"] = torch.zeros(B, H, W, 3, device="cuda")
        out["adj_camera_ray_dir"] = torch.zeros(B, H, W, 3, device="cuda")
    if tf:
        out["adj_tf"] = torch.zeros(scene.tf.shape, device="cuda")
    adj = capi.make_adjoint_outputs(tf_delayed_accumulation=delayed, **out)
    c, t, a = torch.from_numpy(color).cuda(), torch.from_numpy(term).cuda(), torch.from_numpy(adj_color).cuda()
    lib.render_adjoint(inp, c, t, a, adj, torch.cuda.current_stream().cuda_stream, options)
    torch.cuda.synchronize()
    return {k: v.cpu().numpy() for k, v in out.items()}


def forward_gradients(scene, D, *, d_stepsize=-1, d_ray_start=(-1, -1, -1), d_ray_dir=(-1, -1, -1), d_tf=None,
                      lib=None, options=None):
    lib = lib or capi.load()
    B, H, W = scene.B, scene.H, scene.W
    inp = scene.inputs(to_cuda)
    dtf = torch.from_numpy(d_tf).cuda() if d_tf is not None else None
    st = capi.make_fd_settings(D, d_stepsize, d_ray_start, d_ray_dir, dtf, d_tf is not None)
    color = torch.empty(B, H, W, 4, device="cuda")
    term = torch.empty(B, H, W, dtype=torch.int32, device="cuda")
    grad = torch.zeros(B, H, W, D, 4, device="cuda")
    lib.render_forward_gradients(inp, st, color, term, grad, torch.cuda.current_stream().cuda_stream, options)
    torch.cuda.synchronize()
    return color.cpu().numpy(), term.cpu().numpy(), grad.cpu().numpy()


def forward_variables_to_gradients(fv, adj_color, D, *, tf_shape=None, d_stepsize=-1, d_ray_start=(-1, -1, -1),
                                   d_ray_dir=(-1, -1, -1), d_tf=None, lib=None):
    lib = lib or capi.load()
    B, H, W = fv.shape[:3]
    dtf = torch.from_numpy(d_tf).cuda() if d_tf is not None else None
    st = capi.make_fd_settings(D, d_stepsize, d_ray_start, d_ray_dir, dtf, d_tf is not None)
    out = {}
    if d_stepsize >= 0:
        out["adj_step_size"] = torch.zeros(B, H, W, device="cuda")
    if any(v >= 0 for v in tuple(d_ray_start) + tuple(d_ray_dir)):
        out["adj_camera_ray_start"] = torch.zeros(B, H, W, 3, device="cuda")
        out["adj_camera_ray_dir"] = torch.zeros(B, H, W, 3, device="cuda")
    if d_tf is not None:
        out["adj_tf"] = torch.zeros(tf_shape, device="cuda")
    adj = capi.make_adjoint_outputs(**out)
    f, a = torch.from_numpy(fv).cuda(), torch.from_numpy(adj_color).cuda()
    lib.forward_variables_to_gradients(f, a, st, adj, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    return {k: v.cpu().numpy() for k, v in out.items()}
