"""GPU: Camera::generateRays (forward + backward) and Renderer::compareToImage on sm_100a, through the
C-ABI, against the CPU oracle and the committed fixtures of the reference-built shim
(tests/golden/ref_camera_*.npz). Tolerances: rays rel-L2 <= 1e-6 (element-wise arithmetic), sums over
pixels rel-L2 <= 1e-4 (summation order differs)."""
import os

import numpy as np
import pytest
import torch

import camera_cases as cc
import oracle_loader as ol
import scenes
from diffdvr_b200 import capi

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
RAYS = cc.ray_cases()
COMPARE = cc.compare_cases()


def _wrap(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def _unwrap(t):
    return t.cpu().numpy()


def _gpu(fn, case):
    return fn(capi.load(), case, np.float32, _wrap, _unwrap, torch.cuda.synchronize)


@pytest.mark.parametrize("name", sorted(RAYS))
def test_rays_match_oracle_and_golden(name):
    got = _gpu(cc.run_rays, RAYS[name])
    want = cc.run_rays(ol.port(32), RAYS[name])
    gold = np.load(os.path.join(GOLDEN, f"ref_camera_{name}.npz"))
    for ref in (want, gold):
        assert np.array_equal(got["ray_start"], ref["ray_start"])
        assert scenes.rel_l2(got["ray_dir"], ref["ray_dir"]) <= 1e-6
        assert scenes.rel_l2(got["adj_viewport"], ref["adj_viewport"]) <= 1e-4


@pytest.mark.parametrize("name", sorted(COMPARE))
def test_compare_matches_oracle_and_golden(name):
    got = _gpu(cc.run_compare, COMPARE[name])
    want = cc.run_compare(ol.port(32), COMPARE[name])
    gold = np.load(os.path.join(GOLDEN, f"ref_camera_{name}.npz"))
    for ref in (want, gold):
        for k in ("difference", "gradients"):
            assert scenes.rel_l2(got[k], ref[k]) <= 1e-6, k
        for k in ("difference_reduced", "gradients_reduced"):
            assert scenes.rel_l2(got[k], ref[k]) <= 1e-4, k


def test_rays_strided_outputs_equal_packed():
    """channel-first (non-packed) output tensors take the scalar path and give identical rays."""
    lib = capi.load()
    B, H, W, fov = 2, 16, 32, 0.6
    vp = _wrap(cc.make_viewports(B, np.random.RandomState(3)))
    s0, d0 = torch.zeros(B, H, W, 3, device="cuda"), torch.zeros(B, H, W, 3, device="cuda")
    lib.generate_rays(vp, fov, s0, d0)
    s1 = torch.zeros(B, 3, H, W, device="cuda").permute(0, 2, 3, 1)
    d1 = torch.zeros(B, 3, H, W, device="cuda").permute(0, 2, 3, 1)
    lib.generate_rays(vp, fov, s1, d1)
    torch.cuda.synchronize()
    assert torch.equal(s0, s1) and torch.equal(d0, d1)
    g = torch.randn(B, H, W, 3, device="cuda")
    a0, a1 = torch.empty(B, 3, 3, device="cuda"), torch.empty(B, 3, 3, device="cuda")
    lib.generate_rays_adjoint(vp, fov, g, g, a0)
    gp = g.permute(0, 3, 1, 2).contiguous().permute(0, 2, 3, 1)
    lib.generate_rays_adjoint(vp, fov, gp, gp, a1)
    torch.cuda.synchronize()
    assert scenes.rel_l2(a0.cpu().numpy(), a1.cpu().numpy()) <= 1e-5


def test_rays_full_size_against_oracle_and_marcher():
    """C3-sized rays (8 views @512^2): oracle parity on one view, unit directions, adjoint vs the fp64 oracle."""
    lib = capi.load()
    B, H, W, fov = 8, 512, 512, np.radians(45.0)
    vp_np = cc.make_viewports(B, np.random.RandomState(5))
    vp = _wrap(vp_np)
    start, direction = torch.empty(B, H, W, 3, device="cuda"), torch.empty(B, H, W, 3, device="cuda")
    lib.generate_rays(vp, fov, start, direction)
    torch.cuda.synchronize()
    assert torch.allclose(direction.norm(dim=-1), torch.ones(B, H, W, device="cuda"), atol=1e-6)
    ws, wd = np.zeros((1, H, W, 3), np.float32), np.zeros((1, H, W, 3), np.float32)
    ol.port(32).generate_rays(vp_np[3:4], fov, ws, wd)
    assert np.array_equal(start[3:4].cpu().numpy(), ws)
    assert scenes.rel_l2(direction[3:4].cpu().numpy(), wd) <= 1e-6
    # adjoint of a linear functional: <adj, rays(vp)> summed == size-independent check against fp64 oracle
    gs, gd = torch.randn(B, H, W, 3, device="cuda"), torch.randn(B, H, W, 3, device="cuda")
    adj = torch.empty(B, 3, 3, device="cuda")
    lib.generate_rays_adjoint(vp, fov, gs, gd, adj)
    torch.cuda.synchronize()
    want = np.zeros((1, 3, 3), np.float64)
    ol.port(64).generate_rays_adjoint(vp_np[3:4].astype(np.float64), fov, gs[3:4].cpu().numpy().astype(np.float64),
                                      gd[3:4].cpu().numpy().astype(np.float64), want)
    assert scenes.rel_l2(adj[3:4].cpu().numpy(), want) <= 1e-4


def test_compare_reduce_equals_mean_of_unreduced_full_size():
    lib = capi.load()
    B, H, W, D = 8, 512, 512, 6
    g = torch.Generator(device="cuda").manual_seed(1)
    color = torch.rand(B, H, W, 4, device="cuda", generator=g)
    ref = torch.rand(B, H, W, 4, device="cuda", generator=g)
    grads = torch.randn(B, H, W, D, 4, device="cuda", generator=g)
    diff, gout = torch.empty(B, H, W, 1, device="cuda"), torch.empty(B, H, W, D, device="cuda")
    lib.compare_to_image(color, grads, ref, diff, gout, False)
    dr, gr = torch.empty(1, device="cuda"), torch.empty(D, device="cuda")
    lib.compare_to_image(color, grads, ref, dr, gr, True)
    torch.cuda.synchronize()
    v = color - ref
    want = v.norm(dim=-1, keepdim=True)
    assert torch.allclose(diff, want, rtol=1e-6, atol=1e-7)
    wg = (v.unsqueeze(3) * grads).sum(-1) / want
    assert scenes.rel_l2(gout.cpu().numpy(), wg.cpu().numpy()) <= 1e-6
    assert abs(dr.item() - want.double().mean().item()) <= 1e-5 * want.double().mean().item()
    assert scenes.rel_l2(gr.cpu().numpy(), wg.double().mean(dim=(0, 1, 2)).cpu().numpy()) <= 1e-4


def test_argument_errors():
    lib = capi.load()
    vp = torch.zeros(2, 3, 3, device="cuda")
    with pytest.raises(capi.DvrError):
        lib.generate_rays(vp, 0.5, torch.zeros(1, 4, 4, 3, device="cuda"), torch.zeros(2, 4, 4, 3, device="cuda"))
    with pytest.raises(capi.DvrError):
        lib.compare_to_image(torch.zeros(1, 4, 4, 4, device="cuda"), torch.zeros(1, 4, 4, 2, 4, device="cuda"),
                             torch.zeros(1, 4, 4, 4, device="cuda"), torch.zeros(1, device="cuda"),
                             torch.zeros(3, device="cuda"), True)
