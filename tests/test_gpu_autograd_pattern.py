"""GPU: the reference's own training-loop usage of `pyrenderer` -- a torch.autograd.Function around
Renderer.render_forward / render_adjoint, and around render_forward_gradients / forward_variables_to_gradients
(pytests/diffdvr/renderer.py:60-264) -- run against this repository's module. The checks are the ones the reference's
experiments rely on: adjoint and forward-mode give the same parameter gradients, the gradients agree with central
differences of the rendered loss, and an optimiser step through the module reduces the loss."""
import numpy as np
import pytest
import torch

import pyr_runner as pr
import scenes
from diffdvr_b200 import capi

pytestmark = pytest.mark.gpu
pyr = pr.pyr


class RenderFn(torch.autograd.Function):
    """inputs: a pyrenderer.RendererInputs whose tf / volume / camera are replaced per call, as in the reference"""

    @staticmethod
    def forward(ctx, camera, tf, volume, inputs, fov, opt_tf, opt_volume, adjoint_method, delayed):
        inputs.camera = pyr.CameraReferenceFrame(camera, fov)
        inputs.tf = tf
        inputs.volume = volume
        B = max(camera.shape[0], tf.shape[0], volume.shape[0])
        W, H = inputs.screen_size.x, inputs.screen_size.y
        color = torch.empty(B, H, W, 4, device=volume.device)
        term = torch.empty(B, H, W, dtype=torch.int32, device=volume.device)
        outputs = pyr.RendererOutputs(color, term)
        ctx.cfg = (inputs, fov, opt_tf, opt_volume, adjoint_method, delayed)
        if adjoint_method:
            pyr.Renderer.render_forward(inputs, outputs)
            ctx.save_for_backward(camera, tf, volume, color, term)
            ctx.settings = None
        else:
            _, R, C = tf.shape
            s = pyr.ForwardDifferencesSettings()
            s.has_tf_derivatives = True
            s.d_tf = torch.arange(R * C, dtype=torch.int32, device=tf.device).reshape(1, R, C)
            s.D = R * C
            fv = torch.zeros(B, H, W, s.D, 4, device=volume.device)
            pyr.Renderer.render_forward_gradients(inputs, s, outputs, fv)
            ctx.save_for_backward(camera, tf, volume, fv)
            ctx.settings = s
        return color

    @staticmethod
    def backward(ctx, grad_color):
        inputs, fov, opt_tf, opt_volume, adjoint_method, delayed = ctx.cfg
        grad_tf = grad_volume = None
        g = pyr.AdjointOutputs()
        if adjoint_method:
            camera, tf, volume, color, term = ctx.saved_tensors
            inputs.camera = pyr.CameraReferenceFrame(camera, fov)
            inputs.tf = tf
            inputs.volume = volume
            if opt_tf:
                grad_tf = torch.zeros_like(tf)
                g.has_tf_derivatives = True
                g.tf_delayed_accumulation = delayed
                g.adj_tf = grad_tf
            if opt_volume:
                grad_volume = torch.zeros_like(volume)
                g.has_volume_derivatives = True
                g.adj_volume = grad_volume
            pyr.Renderer.render_adjoint(inputs, pyr.RendererOutputs(color, term), grad_color.contiguous(), g)
        else:
            camera, tf, volume, fv = ctx.saved_tensors
            grad_tf = torch.zeros_like(tf)
            g.has_tf_derivatives = True
            g.adj_tf = grad_tf
            pyr.Renderer.forward_variables_to_gradients(fv, grad_color.contiguous(), ctx.settings, g)
        return None, grad_tf, grad_volume, None, None, None, None, None, None


def _setup(n=24, W=32, H=24, R=8, B=2):
    sc = scenes.make_scene(n=n, W=W, H=H, B=B, camera_mode=capi.CAMERA_REFERENCE_FRAME, tf_mode=capi.TF_TEXTURE, R=R,
                           seed=2, opacity=3.0, step_voxels=0.5)
    inputs = pr.make_inputs(sc)
    camera = torch.from_numpy(sc.camera_matrix).cuda()
    tf = torch.from_numpy(sc.tf).cuda()
    volume = torch.from_numpy(np.ascontiguousarray(sc.volume)).cuda()
    target = torch.from_numpy(np.random.RandomState(5).rand(B, H, W, 4).astype(np.float32)).cuda()
    return sc, inputs, camera, tf, volume, target


def _loss(color, target):
    return ((color - target) ** 2).mean()


def test_adjoint_and_forward_mode_give_the_same_tf_gradient():
    sc, inputs, camera, tf, volume, target = _setup()
    grads = {}
    for name, adjoint_method, delayed in (("adjoint", True, False), ("adjoint-delayed", True, True), ("forward", False, False)):
        t = tf.clone().requires_grad_(True)
        color = RenderFn.apply(camera, t, volume, inputs, float(sc.fov_y), True, False, adjoint_method, delayed)
        _loss(color, target).backward()
        grads[name] = t.grad.cpu().numpy()
    assert np.abs(grads["adjoint"]).max() > 0
    assert scenes.rel_l2(grads["adjoint-delayed"], grads["adjoint"]) <= 1e-5
    assert scenes.rel_l2(grads["forward"], grads["adjoint"]) <= 1e-4


def test_gradients_agree_with_central_differences():
    sc, inputs, camera, tf, volume, target = _setup()
    t = tf.clone().requires_grad_(True)
    v = volume.clone().requires_grad_(True)
    color = RenderFn.apply(camera, t, v, inputs, float(sc.fov_y), True, True, True, False)
    _loss(color, target).backward()

    def loss_at(tf_, vol_):
        with torch.no_grad():
            return float(_loss(RenderFn.apply(camera, tf_, vol_, inputs, float(sc.fov_y), False, False, True, False),
                               target).double())

    rng = np.random.RandomState(0)
    # directional derivatives: fp32 rendering allows ~1e-2 relative agreement at this step
    for what, param, grad in (("tf", tf, t.grad), ("volume", volume, v.grad)):
        # a direction along the gradient's signs: the directional derivative is a sum of same-signed terms, so the
        # comparison is not a small difference of large cancelling contributions (a plain random direction on the
        # 24^3 volume leaves |an| ~ 1e-2 of sum |grad_i d_i|, where the kinks of the piecewise-linear TF that the
        # +-eps renders straddle decide the central difference)
        d = torch.from_numpy(rng.uniform(0.5, 1.0, size=tuple(param.shape)).astype(np.float32)).cuda() * torch.sign(grad)
        if what == "tf":
            d[..., :3] *= 0.3
        eps = 2e-3
        args_p = (param + eps * d, volume) if what == "tf" else (tf, param + eps * d)
        args_m = (param - eps * d, volume) if what == "tf" else (tf, param - eps * d)
        fd = (loss_at(*args_p) - loss_at(*args_m)) / (2 * eps)
        an = float((grad.double() * d.double()).sum())
        assert abs(fd - an) <= 2e-2 * max(abs(fd), abs(an)) + 1e-7, (what, fd, an)


def test_an_optimiser_step_through_the_module_reduces_the_loss():
    sc, inputs, camera, tf, volume, _ = _setup()
    with torch.no_grad():
        target = RenderFn.apply(camera, tf, volume, inputs, float(sc.fov_y), False, False, True, False).clone()
    v = (volume * 0.5 + 0.1).clone().requires_grad_(True)
    opt = torch.optim.Adam([v], lr=0.02)
    losses = []
    n0 = pyr.launch_count()
    for _ in range(8):
        opt.zero_grad()
        loss = _loss(RenderFn.apply(camera, tf, v, inputs, float(sc.fov_y), False, True, True, False), target)
        loss.backward()
        opt.step()
        with torch.no_grad():
            v.clamp_(0.0, 1.0)
        losses.append(float(loss))
    assert pyr.launch_count() - n0 >= 16
    assert losses[-1] < 0.7 * losses[0], losses
