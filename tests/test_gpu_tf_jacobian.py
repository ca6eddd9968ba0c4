"""GPU: forward-mode with TRANSFER-FUNCTION tangents in one marching pass (diffdvr_b200/csrc/dvr_tfjac.cuh) against
the slot-major kernel it replaces (the reference's scheme, cudAD::fvar<real_t, D> in chunks of 8 slots: ceil(D / 8)
marches) and against the oracle's restated dual numbers. The reference seeds these tangents at
renderer/renderer_tf.cuh:121-154 (Identity), 255-288 (Texture), 436-485 (Linear), 521-560 (Gaussian)."""
import numpy as np
import pytest
import torch

import gpu_runner as gr
import oracle_loader as ol
import scenes
from diffdvr_b200 import capi
from test_gpu_forward_mode import oracle_forward_gradients

pytestmark = pytest.mark.gpu

SLOT_MAJOR = dict(forward_mode_tf=1)


def _tf_slots(sc, first=0):
    R, C = sc.tf.shape[1:]
    return np.arange(first, first + R * C, dtype=np.int32).reshape(1, R, C)


@pytest.mark.parametrize("tfm,R", [(capi.TF_IDENTITY, 1), (capi.TF_TEXTURE, 64), (capi.TF_TEXTURE, 2), (capi.TF_LINEAR, 7),
                                   (capi.TF_GAUSSIAN, 3), (capi.TF_GAUSSIAN, 9)])
@pytest.mark.parametrize("blend", [capi.BLEND_BEER_LAMBERT, capi.BLEND_ALPHA])
@pytest.mark.parametrize("filt", [capi.FILTER_NEAREST, capi.FILTER_TRILINEAR])
def test_one_pass_equals_the_slot_major_passes(tfm, R, blend, filt):
    sc = scenes.make_scene(n=28, W=40, H=24, B=2, camera_mode=capi.CAMERA_REFERENCE_FRAME, tf_mode=tfm, R=R,
                           blend_mode=blend, filter_mode=filt, seed=21, opacity=4.0, step_voxels=0.4)
    d_tf = _tf_slots(sc)
    D = d_tf.size
    lib = capi.load()
    n0 = lib.launch_count()
    c1, t1, g1 = gr.forward_gradients(sc, D, d_tf=d_tf, lib=lib)
    one_pass = lib.launch_count() - n0
    n0 = lib.launch_count()
    c8, t8, g8 = gr.forward_gradients(sc, D, d_tf=d_tf, lib=lib, options=capi.make_call_options(**SLOT_MAJOR))
    slot_major = lib.launch_count() - n0
    # forward + Jacobian (+ the QUAD copy; the zero-fill of a dense output is a memset, not a launch)
    assert one_pass <= 3, one_pass
    assert slot_major >= (D + 7) // 8
    assert np.array_equal(t1, t8) and scenes.rel_l2(c1, c8) <= 1e-6
    assert np.abs(g8).max() > 0
    assert scenes.rel_l2(g1, g8) <= 2e-5, scenes.rel_l2(g1, g8)
    # every slot on its own (a relative bound on the whole Jacobian would hide a wrong small column)
    for d in range(D):
        ref = np.abs(g8[..., d, :]).max()
        if ref > 1e-6 * np.abs(g8).max():
            assert scenes.rel_l2(g1[..., d, :], g8[..., d, :]) <= 2e-4, d


@pytest.mark.parametrize("tfm,R", [(capi.TF_TEXTURE, 6), (capi.TF_LINEAR, 4), (capi.TF_IDENTITY, 1), (capi.TF_GAUSSIAN, 4)])
def test_one_pass_against_the_restated_dual_numbers(tfm, R):
    """The oracle pushes fvar<D> through the reference's kernel; D = 24 / 20 / 2 here, more than one chunk of the
    slot-major kernel."""
    sc = scenes.make_scene(n=24, W=32, H=20, B=2, camera_mode=capi.CAMERA_RAY_START_DIR, tf_mode=tfm, R=R, seed=8,
                           opacity=3.0)
    d_tf = _tf_slots(sc)
    D = d_tf.size
    want_c, want_t, want_g = oracle_forward_gradients(ol.port(32), sc, D, d_tf=d_tf)
    got_c, got_t, got_g = gr.forward_gradients(sc, D, d_tf=d_tf)
    assert scenes.rel_l2(got_c, want_c) <= 1e-5 and (got_t != want_t).mean() <= 1e-3
    same = (got_t == want_t)[..., None, None]
    assert scenes.rel_l2(got_g * same, want_g * same) <= 1e-4


@pytest.mark.parametrize("tfm,R", [(capi.TF_TEXTURE, 5), (capi.TF_LINEAR, 4), (capi.TF_GAUSSIAN, 3)])
def test_slot_maps_shared_missing_and_mixed_with_ray_tangents(tfm, R):
    """Slots shared between TF entries and with a ray input add up; entries without a slot (-1) are dropped; the
    ray / stepsize tangents of the same call come from the input-major kernel and are kept."""
    sc = scenes.make_scene(n=22, W=24, H=16, B=2, camera_mode=capi.CAMERA_RAY_START_DIR, tf_mode=tfm, R=R, seed=14,
                           opacity=3.0)
    C = sc.tf.shape[2]
    d_tf = (3 + np.arange(R * C, dtype=np.int32) % 7).reshape(1, R, C)   # 7 TF slots, each shared by several entries
    d_tf[0, 1, 3] = -1
    d_tf[0, 0, 0] = 0                                                     # shared with ray start x
    D = 12                                                                # slots 10, 11 stay unused
    kw = dict(d_ray_start=(0, 1, 2), d_stepsize=9, d_tf=d_tf)
    want_c, want_t, want_g = oracle_forward_gradients(ol.port(32), sc, D, **dict(kw))
    got_c, got_t, got_g = gr.forward_gradients(sc, D, **kw)
    _, _, g8 = gr.forward_gradients(sc, D, options=capi.make_call_options(**SLOT_MAJOR), **kw)
    assert scenes.rel_l2(got_c, want_c) <= 1e-5
    assert not got_g[..., 10:, :].any()
    assert scenes.rel_l2(got_g[..., 3:9, :], g8[..., 3:9, :]) <= 5e-5   # pure TF slots: the two product kernels agree
    if scenes.rel_l2(got_g, want_g) > 1e-4:
        _, _, g64 = oracle_forward_gradients(ol.port(64), sc, D, np.float64, **dict(kw))
        scenes.assert_position_gradient_close("jacobian", got_g, want_g, g64, 1e-4,
                                              scenes.face_sensitive_rays(sc, want_t))


def test_strided_jacobian_output():
    """A gradient tensor that is a permuted view (slot-major in memory): the zero-fill kernel and the scalar
    read-modify-writes instead of cudaMemsetAsync and 16-byte accesses."""
    sc = scenes.make_scene(n=20, W=24, H=16, B=2, camera_mode=capi.CAMERA_REFERENCE_FRAME, tf_mode=capi.TF_TEXTURE, R=4,
                           seed=3, opacity=3.0)
    d_tf = _tf_slots(sc)
    D = d_tf.size
    _, _, want = gr.forward_gradients(sc, D, d_tf=d_tf)
    lib = capi.load()
    B, H, W = sc.B, sc.H, sc.W
    inp = sc.inputs(gr.to_cuda)
    dtf = torch.from_numpy(d_tf).cuda()
    st = capi.make_fd_settings(D, d_tf=dtf, has_tf_derivatives=True)
    color = torch.empty(B, H, W, 4, device="cuda")
    term = torch.empty(B, H, W, dtype=torch.int32, device="cuda")
    store = torch.full((B, D, 4, H, W + 3), 7.0, device="cuda")       # padded rows: the padding must stay untouched
    grad = store[..., :W].permute(0, 3, 4, 1, 2)
    lib.render_forward_gradients(inp, st, color, term, grad, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    assert np.array_equal(grad.cpu().numpy(), want)
    assert bool((store[..., W:] == 7.0).all())


def test_gaussian_tables_too_large_for_the_moment_table_take_the_slot_major_passes():
    """15 moments per row and thread in shared memory: R = 40 does not fit, the call falls back to ceil(D / 8) passes
    and still matches the forced slot-major result bit for bit (it IS that path)."""
    sc = scenes.make_scene(n=16, W=16, H=8, B=1, camera_mode=capi.CAMERA_REFERENCE_FRAME, tf_mode=capi.TF_GAUSSIAN, R=40,
                           seed=5, opacity=2.0)
    d_tf = _tf_slots(sc)
    D = d_tf.size
    lib = capi.load()
    n0 = lib.launch_count()
    _, _, g = gr.forward_gradients(sc, D, d_tf=d_tf, lib=lib)
    assert lib.launch_count() - n0 >= (D + 7) // 8
    _, _, g8 = gr.forward_gradients(sc, D, d_tf=d_tf, lib=lib, options=capi.make_call_options(**SLOT_MAJOR))
    assert np.array_equal(g, g8)
