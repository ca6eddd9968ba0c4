"""The seeded scenes behind the golden fixtures tests/golden/ref_*.npz (see golden/make_golden.py)
and the one routine that runs ANY library exporting the C-ABI on them (the reference's compiled host
kernels when generating, the CPU port and the CUDA product when checking)."""
from __future__ import annotations

import numpy as np

import scenes
from diffdvr_b200 import capi

ADJ_SEED = 11


def cases():
    """(name, scene, adjoint flags, forward-mode settings or None)"""
    F, C, T, B = capi.FILTER_TRILINEAR, capi.CAMERA_REFERENCE_FRAME, capi.TF_TEXTURE, capi.BLEND_BEER_LAMBERT
    mk = scenes.make_scene
    all_flags = dict(volume=True, step=True, camera=True, tf=True)
    cam6 = dict(D=6, d_ray_start=(0, 1, 2), d_ray_dir=(3, 4, 5))
    out = [
        ("tri_frame_texture_beer", mk(n=16, W=14, H=10, B=2, camera_mode=C, tf_mode=T, R=8, seed=21, opacity=4.0),
         all_flags, None),
        ("tri_rays_linear_alpha", mk(n=16, W=12, H=10, B=1, camera_mode=capi.CAMERA_RAY_START_DIR,
                                     tf_mode=capi.TF_LINEAR, R=5, blend_mode=capi.BLEND_ALPHA, seed=22, opacity=4.0),
         all_flags, dict(D=7, d_stepsize=6, **{k: v for k, v in cam6.items() if k != "D"})),
        ("tri_invview_gaussian_beer", mk(n=16, W=12, H=10, B=2, camera_mode=capi.CAMERA_INVERSE_VIEW_MATRIX,
                                         tf_mode=capi.TF_GAUSSIAN, R=3, seed=23, opacity=4.0),
         all_flags, "tf"),
        ("tri_frame_identity_beer", mk(n=16, W=14, H=10, B=1, camera_mode=C, tf_mode=capi.TF_IDENTITY, seed=24,
                                       opacity=3.0), all_flags, "tf"),
        ("near_frame_texture_alpha", mk(n=16, W=12, H=10, B=1, camera_mode=C, tf_mode=T, R=6,
                                        blend_mode=capi.BLEND_ALPHA, filter_mode=capi.FILTER_NEAREST, seed=25,
                                        opacity=4.0), dict(volume=True, step=False, camera=False, tf=True), None),
        ("tri_frame_gaussianlog_beer", mk(n=16, W=12, H=10, B=1, camera_mode=C, tf_mode=capi.TF_GAUSSIAN_LOG, R=3,
                                          seed=26, opacity=1.0), dict(volume=True, step=False, camera=False, tf=True),
         None),
        ("tri_xfast_randvol_texture", mk(n=12, W=11, H=9, B=2, camera_mode=C, tf_mode=T, R=4, seed=27, opacity=4.0,
                                         volume_kind="rand", x_fastest=True, tensor_scalars=True),
         all_flags, "tf"),
        ("pre_frame_beer", mk(n=12, W=12, H=10, B=2, camera_mode=C, filter_mode=capi.FILTER_PRESHADED, seed=28,
                              opacity=4.0), dict(volume=True, step=True, camera=True, tf=False), None),
        ("pre_rays_alpha", mk(n=12, W=11, H=9, B=1, camera_mode=capi.CAMERA_RAY_START_DIR,
                              filter_mode=capi.FILTER_PRESHADED, blend_mode=capi.BLEND_ALPHA, seed=29, opacity=4.0),
         dict(volume=True, step=True, camera=True, tf=False), None),
        ("consistency_tf_texture8", scenes.consistency_tf_scene(T, 8, seed=8, W=24, H=24),
         dict(volume=False, step=False, camera=False, tf=True), "tf"),
        ("consistency_camera_linear4", scenes.consistency_camera_scene(capi.TF_LINEAR, 4, seed=4),
         dict(volume=False, step=False, camera=True, tf=False), cam6),
    ]
    return out


def fg_settings(sc, fg):
    if fg is None:
        return None
    if fg == "tf":
        Rn, C = sc.tf.shape[1:]
        return dict(D=Rn * C, d_tf=np.arange(Rn * C, dtype=np.int32).reshape(1, Rn, C))
    return dict(fg)


def run(lib, sc, flags, fg, runner=None, forward_from=None, runner_kwargs=None):
    """Runs forward, adjoint and, if ``fg``, forward-mode on ``sc`` through ``lib`` (host pointers) or
    through ``runner`` (the GPU runner module, device pointers). Returns a dict of numpy arrays."""
    rk = runner_kwargs or {}
    if runner is None:
        color, term = scenes.np_forward(lib, sc)
    else:
        color, term = runner.forward(sc, **rk)
    out = {"color": color, "term": term}
    adj_color = np.random.RandomState(ADJ_SEED).randn(*color.shape).astype(np.float32)
    # the adjoint consumes the GOLDEN forward outputs when given, so that a +-1 terminationIndex
    # flip of a borderline ray cannot leak into the gradient comparison
    c_in, t_in = forward_from if forward_from is not None else (color, term)
    if runner is None:
        adj = scenes.np_adjoint(lib, sc, c_in, t_in, adj_color, **flags)
    else:
        adj = runner.adjoint(sc, c_in, t_in, adj_color, **flags, **rk)
    out.update(adj)
    st = fg_settings(sc, fg)
    if st is not None:
        D = st.pop("D")
        if runner is None:
            B, H, W = sc.B, sc.H, sc.W
            inp = sc.inputs()
            s = capi.make_fd_settings(D, d_tf=st.get("d_tf"), has_tf_derivatives="d_tf" in st,
                                      **{k: v for k, v in st.items() if k != "d_tf"})
            c2 = np.zeros((B, H, W, 4), np.float32)
            t2 = np.zeros((B, H, W), np.int32)
            g = np.zeros((B, H, W, D, 4), np.float32)
            lib.render_forward_gradients(inp, s, c2, t2, g)
        else:
            c2, t2, g = runner.forward_gradients(sc, D, **st, **rk)
        out["fg_color"], out["fg_term"], out["fg_gradients"] = c2, t2, g
    return out
