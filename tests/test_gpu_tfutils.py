"""GPU: TFUtils::preshadeVolume and its density adjoint on sm_100a through the C-ABI against the CPU oracle and the
fixtures of the reference-built shim (tests/golden/ref_preshade_*.npz); element-wise arithmetic, tolerance
rel-L2 <= 1e-5 on colours and <= 1e-4 on gradients. Plus: the pre-shaded marcher consumes the output, and a
full-size (256^3) run against properties that do not need the oracle."""
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
CASES = cc.preshade_cases()


def _wrap(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


@pytest.mark.parametrize("name", sorted(CASES))
def test_preshade_matches_oracle_and_golden(name):
    got = cc.run_preshade(capi.load(), CASES[name], np.float32, _wrap, lambda t: t.cpu().numpy(),
                          torch.cuda.synchronize)
    want = cc.run_preshade(ol.port(32), CASES[name])
    gold = np.load(os.path.join(GOLDEN, f"ref_{name}.npz"))
    for ref in (want, gold):
        assert scenes.rel_l2(got["color"], ref["color"]) <= 1e-5
        assert scenes.rel_l2(got["adj_volume"], ref["adj_volume"]) <= 1e-4


def test_strided_tensors():
    """x-fastest density volume and channel-last colour volume (non-contiguous views)."""
    lib = capi.load()
    rng = np.random.RandomState(4)
    tf = _wrap(scenes.random_tf(capi.TF_LINEAR, 5, rng, opacity=4.0))
    vol = torch.rand(1, 9, 7, 5, device="cuda")
    c0 = torch.empty(4, 9, 7, 5, device="cuda")
    lib.preshade_volume(vol, tf, capi.TF_LINEAR, c0)
    vol_t = vol.permute(0, 3, 2, 1).contiguous().permute(0, 3, 2, 1)
    c1 = torch.empty(9, 7, 5, 4, device="cuda").permute(3, 0, 1, 2)
    lib.preshade_volume(vol_t, tf, capi.TF_LINEAR, c1)
    torch.cuda.synchronize()
    assert not vol_t.is_contiguous() and not c1.is_contiguous()
    assert torch.equal(c0, c1)


def test_full_size_identity_and_marcher_consumes_output():
    lib = capi.load()
    n = 256
    vol = torch.rand(1, n, n, n, device="cuda") * 1.2 - 0.1
    tf = torch.tensor([[[3.0, 0.7]]], device="cuda")
    color = torch.empty(4, n, n, n, device="cuda")
    lib.preshade_volume(vol, tf, capi.TF_IDENTITY, color)
    dc = vol[0].clamp(0, 1)
    assert torch.equal(color[0], dc * 0.7) and torch.equal(color[3], dc * 3.0)
    adj_c = torch.randn(4, n, n, n, device="cuda")
    adj_v = torch.empty(1, n, n, n, device="cuda")
    lib.preshade_volume_adjoint(vol, tf, capi.TF_IDENTITY, adj_c, adj_v)
    inside = (vol[0] > 0) & (vol[0] < 1)
    want = torch.where(inside, (adj_c[0] * 0.7 + adj_c[1] * 0.7 + adj_c[2] * 0.7) + adj_c[3] * 3.0,
                       torch.zeros((), device="cuda"))
    assert torch.allclose(adj_v[0], want, rtol=1e-6, atol=1e-6)
    # the pre-shaded marcher renders it
    vp = _wrap(cc.make_viewports(1, np.random.RandomState(1)))
    inp = capi.Inputs(screen_size=(64, 48), volume=color, box_min=(-0.5, -0.5, -0.5), box_size=(1, 1, 1),
                      camera_mode=capi.CAMERA_REFERENCE_FRAME, camera_matrix=vp, camera_fov_y_radians=0.7,
                      step_size=0.5 / n, tf_mode=capi.TF_PRESHADED, tf=tf, blend_mode=capi.BLEND_BEER_LAMBERT,
                      volume_filter_mode=capi.FILTER_PRESHADED)
    img = torch.empty(1, 48, 64, 4, device="cuda")
    term = torch.empty(1, 48, 64, dtype=torch.int32, device="cuda")
    lib.render_forward(inp, img, term)
    torch.cuda.synchronize()
    assert img[..., 3].max().item() > 0.5 and torch.isfinite(img).all()


def test_argument_errors():
    lib = capi.load()
    vol = torch.zeros(1, 4, 4, 4, device="cuda")
    with pytest.raises(capi.DvrError):  # Texture TF needs 4 channels
        lib.preshade_volume(vol, torch.zeros(1, 4, 5, device="cuda"), capi.TF_TEXTURE, torch.zeros(4, 4, 4, 4, device="cuda"))
    with pytest.raises(capi.DvrError):  # colour volume must be 4*X*Y*Z
        lib.preshade_volume(vol, torch.zeros(1, 4, 4, device="cuda"), capi.TF_TEXTURE, torch.zeros(3, 4, 4, 4, device="cuda"))
