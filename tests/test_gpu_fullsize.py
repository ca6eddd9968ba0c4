"""GPU: the marcher at BASELINE.json's full size (C4: 256^3 volume, 512^2 views, step 0.2 voxel, Identity TF,
Beer-Lambert, ReferenceFrame camera) through properties that need no oracle -- the oracle takes minutes per view at this
size; it checks the same code paths on small scenes in tests/test_gpu_parity.py:
  * determinism of the forward render and view-batch independence (bit-exact),
  * both volume layouts give the same image bit for bit and the same gradient to summation order,
  * the adjoint is LINEAR in adj_color,
  * duality <adj_color, dI/dv . dv> == <adj_volume, dv> against a central difference of the render along a smooth
    direction dv (the defining property of the adjoint),
  * an empty volume renders nothing and a zero adjoint colour produces zero gradients,
  * the texture-TF / stepsize / TF-gradient outputs of C5's flag set are linear too."""
import math

import numpy as np
import pytest
import torch

from diffdvr_b200 import capi, synthetic

pytestmark = pytest.mark.gpu
N, RES, VIEWS = 256, 512, 2


@pytest.fixture(scope="module")
def world():
    lib = capi.load()
    dev = torch.device("cuda", 0)
    volume = (0.5 * synthetic.skull(N) + 0.05).to(dev)  # densities in [0.05, 0.55]: no clamp of the Identity TF
    lat, lon = synthetic.fibonacci_sphere(16)  # two of the bench's camera positions
    vps = synthetic.viewport_from_sphere((0, 0, 0), lon[[3, 11]], lat[[3, 11]], 1.4641, "Ym")
    vps = torch.from_numpy(np.ascontiguousarray(vps, np.float32)).to(dev)
    box_min, box_size = synthetic.default_box(N)

    def inputs(vol, vp=vps, tf_mode=capi.TF_IDENTITY, tf=None, step=0.2 / N):
        tf = tf if tf is not None else torch.from_numpy(synthetic.tf_identity()).to(dev)
        return capi.Inputs(screen_size=(RES, RES), volume=vol, box_min=box_min, box_size=box_size,
                           camera_mode=capi.CAMERA_REFERENCE_FRAME, camera_matrix=vp,
                           camera_fov_y_radians=math.radians(45.0), step_size=step, tf_mode=tf_mode, tf=tf,
                           blend_mode=capi.BLEND_BEER_LAMBERT)

    def forward(inp, B=VIEWS, options=None):
        color = torch.empty(B, RES, RES, 4, device=dev)
        term = torch.empty(B, RES, RES, dtype=torch.int32, device=dev)
        lib.render_forward(inp, color, term, options=options)
        torch.cuda.synchronize()
        return color, term

    def adjoint(inp, color, term, adj_color, options=None, **extra):
        adj_volume = torch.zeros(1, N, N, N, device=dev)
        adj = capi.make_adjoint_outputs(adj_volume=adj_volume, **extra)
        lib.render_adjoint(inp, color, term, adj_color, adj, options=options)
        torch.cuda.synchronize()
        return adj_volume

    return dict(lib=lib, dev=dev, volume=volume, vps=vps, inputs=inputs, forward=forward, adjoint=adjoint)


def rel(a, b):
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


def test_forward_is_deterministic_and_views_are_independent(world):
    c0, t0 = world["forward"](world["inputs"](world["volume"]))
    c1, t1 = world["forward"](world["inputs"](world["volume"]))
    assert torch.equal(c0, c1) and torch.equal(t0, t1)
    for k in range(VIEWS):  # one view per launch == the same view inside a batch
        ck, tk = world["forward"](world["inputs"](world["volume"], world["vps"][k:k + 1].contiguous()), B=1)
        assert torch.equal(ck[0], c0[k]) and torch.equal(tk[0], t0[k])
    alpha = c0[..., 3]
    assert 0.3 < float((alpha > 0).float().mean()) < 1.0 and float(alpha.max()) <= 1.0
    assert int(t0.max()) > 1000  # 256 voxels * sqrt(3) / 0.2 voxel per step: long rays are really marched


def test_layouts_agree_at_full_size(world):
    lib = world["lib"]
    inp = world["inputs"](world["volume"])
    g = torch.Generator(device="cuda").manual_seed(5)
    out = {}
    for layout in (0, 1):
        opts = capi.make_call_options(volume_layout=layout)
        color, term = world["forward"](inp, options=opts)
        adj_color = torch.randn(color.shape, device="cuda", generator=g) if layout == 0 else adj_color
        out[layout] = (color, term, world["adjoint"](inp, color, term, adj_color, options=opts))
    assert torch.equal(out[0][0], out[1][0]) and torch.equal(out[0][1], out[1][1])
    assert rel(out[1][2], out[0][2]) <= 2e-6


def test_adjoint_is_linear_in_adj_color(world):
    inp = world["inputs"](world["volume"])
    color, term = world["forward"](inp)
    g = torch.Generator(device="cuda").manual_seed(6)
    g1 = torch.randn(color.shape, device="cuda", generator=g)
    g2 = torch.randn(color.shape, device="cuda", generator=g)
    a1, a2 = world["adjoint"](inp, color, term, g1), world["adjoint"](inp, color, term, g2)
    a12 = world["adjoint"](inp, color, term, 0.75 * g1 - 2.0 * g2)
    assert rel(a12, 0.75 * a1 - 2.0 * a2) <= 1e-4
    zero = world["adjoint"](inp, color, term, torch.zeros_like(color))
    assert float(zero.abs().max()) == 0.0


def test_adjoint_is_the_transpose_of_the_directional_derivative(world):
    """<adj_color, (I(v + e dv) - I(v - e dv)) / 2e> == <adjoint(adj_color), dv>; Identity TF + Beer-Lambert is smooth in
    the density, the truncation error of the central difference is O(e^2)."""
    vol = world["volume"]
    ax = torch.linspace(-1, 1, N, device="cuda")
    x, y, z = torch.meshgrid(ax, ax, ax, indexing="ij")
    dv = (torch.sin(3 * x + 1) * torch.cos(2 * y) * torch.sin(2.5 * z + 0.5))[None].contiguous()
    inp = world["inputs"](vol)
    color, term = world["forward"](inp)
    # a smooth adjoint colour keeps the fp32 round-off of the image difference (~1e-7 / e per pixel) averaged out
    adj_color = torch.zeros_like(color)
    adj_color[..., 3] = 1.0 / color[..., 3].numel()
    adj_color[..., 0] = -0.5 / color[..., 3].numel()
    lhs_vec = world["adjoint"](inp, color, term, adj_color)
    rhs = float((lhs_vec.double() * dv.double()).sum())
    eps = 2e-2
    cp, _ = world["forward"](world["inputs"](vol + eps * dv))
    cm, _ = world["forward"](world["inputs"](vol - eps * dv))
    lhs = float((((cp.double() - cm.double()) / (2 * eps)) * adj_color.double()).sum())
    assert abs(lhs - rhs) <= 2e-3 * max(abs(lhs), abs(rhs)), (lhs, rhs)
    assert abs(rhs) > 1e-6


def test_empty_volume_and_tiny_images(world):
    empty = torch.zeros(1, N, N, N, device="cuda")
    color, term = world["forward"](world["inputs"](empty))
    assert float(color.abs().max()) == 0.0
    hit = term > 0
    assert 0.3 < float(hit.float().mean()) < 1.0  # rays through the box are marched to its far side, the others not at all
    # 1 x 1 and ragged images of the same scene (partial 16 x 8 tiles)
    lib = world["lib"]
    for W, H in ((1, 1), (17, 9), (33, 5)):
        inp = world["inputs"](world["volume"])
        inp.screen_size = (W, H)
        c = torch.full((VIEWS, H, W, 4), 7.0, device="cuda")
        t = torch.full((VIEWS, H, W), -1, dtype=torch.int32, device="cuda")
        lib.render_forward(inp, c, t)
        torch.cuda.synchronize()
        assert float(c.max()) <= 1.0 and int(t.min()) >= 0  # every pixel written


def test_c5_flag_set_is_linear_at_full_size(world):
    """adjoint(hasVolume + hasTF + hasStepsize) with the Texture TF (C5's flags) on the 256^3 scene: all three outputs
    are linear in adj_color (shared-memory TF-gradient tables, per-ray stepsize gradients, QUAD gradient scratch)."""
    lib, dev = world["lib"], world["dev"]
    tf = torch.from_numpy(synthetic.tf_texture_from_linear(synthetic.tf_linear_two_peaks(), 64)).to(dev)
    inp = world["inputs"](world["volume"], tf_mode=capi.TF_TEXTURE, tf=tf, step=0.5 / N)
    color, term = world["forward"](inp)
    g = torch.Generator(device="cuda").manual_seed(8)

    def run(adj_color):
        adj_volume = torch.zeros(1, N, N, N, device=dev)
        adj_tf = torch.zeros_like(tf)
        adj_step = torch.zeros(VIEWS, RES, RES, device=dev)
        adj = capi.make_adjoint_outputs(adj_volume=adj_volume, adj_tf=adj_tf, adj_step_size=adj_step)
        lib.render_adjoint(inp, color, term, adj_color, adj)
        torch.cuda.synchronize()
        return adj_volume, adj_tf, adj_step

    g1 = torch.randn(color.shape, device="cuda", generator=g)
    g2 = torch.randn(color.shape, device="cuda", generator=g)
    r1, r2, r12 = run(g1), run(g2), run(g1 + 0.5 * g2)
    for a, b, c, name in zip(r1, r2, r12, ("adj_volume", "adj_tf", "adj_step_size")):
        assert float(c.abs().max()) > 0, name
        assert rel(c, a + 0.5 * b) <= 1e-4, name
