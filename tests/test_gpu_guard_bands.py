"""GPU: out-of-bounds writes, checked without compute-sanitizer (the tool is closed on this GPU pool, see
profiles/r02_sanitizer.md): every output of the three passes lives inside a larger allocation whose margins hold a
sentinel, the kernels run on scenes whose image is ragged against the 16x8 CTA tile and whose volume has three
different extents, and the margins must come back untouched. Repeated runs must also agree bit for bit wherever the
summation order is fixed (everything except the atomically accumulated adj_volume / adj_tf, which must agree to
rounding): an unsynchronised shared-memory table would show up as run-to-run noise far above that."""
import numpy as np
import pytest
import torch

import gpu_runner as gr
import scenes
from diffdvr_b200 import capi

pytestmark = pytest.mark.gpu

SENTINEL = 12345.0  # integral: the int32 buffers hold it exactly
PAD = 64  # elements on both sides


class Guarded:
    """A tensor of `shape` carved out of the middle of a sentinel-filled 1-D allocation."""

    def __init__(self, shape, dtype=torch.float32, fill=0):
        n = int(np.prod(shape))
        self.store = torch.full((n + 2 * PAD,), SENTINEL, device="cuda").to(dtype)
        self.view = self.store[PAD:PAD + n].view(*shape)
        self.view.fill_(fill)
        self.n = n

    def intact(self):
        s = self.store.float()
        return bool((s[:PAD] == SENTINEL).all()) and bool((s[PAD + self.n:] == SENTINEL).all())


def _scene(tf_mode, filt, camera_mode=capi.CAMERA_RAY_START_DIR, R=9):
    # image 37 x 13: ragged against the 16 x 8 CTA tile and the 8 x 4 warp tile; volume 17 x 22 x 19 in the same box
    sc = scenes.make_scene(n=19, W=37, H=13, B=2, camera_mode=camera_mode, tf_mode=tf_mode, R=R, seed=31,
                           opacity=4.0, filter_mode=filt)
    sc.volume = np.random.RandomState(7).rand(1, 17, 22, 19).astype(np.float32)
    return sc


@pytest.mark.parametrize("tf_mode", [capi.TF_IDENTITY, capi.TF_TEXTURE, capi.TF_LINEAR, capi.TF_GAUSSIAN])
@pytest.mark.parametrize("filt", [capi.FILTER_NEAREST, capi.FILTER_TRILINEAR])
@pytest.mark.parametrize("layout", [0, 1])
def test_no_pass_writes_outside_its_outputs(tf_mode, filt, layout):
    sc = _scene(tf_mode, filt)
    lib = capi.load()
    B, H, W = sc.B, sc.H, sc.W
    inp = sc.inputs(gr.to_cuda)
    opts = capi.make_call_options(volume_layout=layout)
    stream = torch.cuda.current_stream().cuda_stream
    color, term = Guarded((B, H, W, 4)), Guarded((B, H, W), torch.int32)
    lib.render_forward(inp, color.view, term.view, stream, opts)
    torch.cuda.synchronize()
    assert color.intact() and term.intact()
    want_c, want_t = gr.forward(sc, lib, options=opts)
    assert np.array_equal(color.view.cpu().numpy(), want_c) and np.array_equal(term.view.cpu().numpy(), want_t)

    adj_color = torch.from_numpy(np.random.RandomState(2).randn(B, H, W, 4).astype(np.float32)).cuda()
    outs = dict(adj_volume=Guarded(tuple(inp.volume.shape)), adj_tf=Guarded(tuple(sc.tf.shape)),
                adj_step_size=Guarded((B, H, W)), adj_camera_ray_start=Guarded((B, H, W, 3)),
                adj_camera_ray_dir=Guarded((B, H, W, 3)))
    adj = capi.make_adjoint_outputs(**{k: v.view for k, v in outs.items()})
    lib.render_adjoint(inp, color.view, term.view, adj_color, adj, stream, opts)
    torch.cuda.synchronize()
    for k, v in outs.items():
        assert v.intact(), k
        assert torch.isfinite(v.view).all(), k
    first = {k: v.view.clone() for k, v in outs.items()}
    for v in outs.values():
        v.view.zero_()
    lib.render_adjoint(inp, color.view, term.view, adj_color, adj, stream, opts)
    torch.cuda.synchronize()
    for k, v in outs.items():
        if k in ("adj_volume", "adj_tf"):   # atomics: order-dependent rounding only
            assert scenes.rel_l2(v.view.cpu().numpy(), first[k].cpu().numpy()) <= 2e-6, k
        else:
            assert torch.equal(v.view, first[k]), k

    if True:
        Rn, C = sc.tf.shape[1:]
        D = 7 + Rn * C
        d_tf = torch.arange(7, D, dtype=torch.int32, device="cuda").reshape(1, Rn, C)
        st = capi.make_fd_settings(D, 6, (0, 1, 2), (3, 4, 5), d_tf, True)
        jac = Guarded((B, H, W, D, 4), fill=float("nan"))
        color2, term2 = Guarded((B, H, W, 4)), Guarded((B, H, W), torch.int32)
        lib.render_forward_gradients(inp, st, color2.view, term2.view, jac.view, stream, opts)
        torch.cuda.synchronize()
        assert jac.intact() and color2.intact() and term2.intact()
        assert torch.isfinite(jac.view).all()   # every slot of every pixel was written (the fill was NaN)
        again = jac.view.clone()
        jac.view.fill_(float("nan"))
        lib.render_forward_gradients(inp, st, color2.view, term2.view, jac.view, stream, opts)
        torch.cuda.synchronize()
        assert torch.equal(jac.view, again)     # no atomics on this path: bit-identical
