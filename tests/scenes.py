"""Seeded test scenes shared by the oracle tests (CPU) and the CUDA parity tests (GPU).

A ``Scene`` keeps every input as a numpy fp32 array / python scalar and can materialise
``capi.Inputs`` for any consumer through a ``conv`` callback (identity for the fp32 CPU oracle,
``astype(float64)`` for the fp64 oracle, ``torch.from_numpy(..).cuda()`` for the product).
Scene definitions follow the reference's own test fixtures where it has them:
unittests/testConsistencyTF.cpp:48-114 and unittests/testConsistencyCamera.cpp:48-118.
"""
from __future__ import annotations

import math
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from diffdvr_b200 import capi, synthetic  # noqa: E402

TF_CHANNELS = {capi.TF_IDENTITY: 2, capi.TF_TEXTURE: 4, capi.TF_LINEAR: 5, capi.TF_GAUSSIAN: 6,
               capi.TF_GAUSSIAN_LOG: 6}


class Scene:
    def __init__(self, *, W, H, volume, box_min, box_size, camera_mode, step_size, tf_mode, tf,
                 blend_mode=capi.BLEND_BEER_LAMBERT, filter_mode=capi.FILTER_TRILINEAR,
                 ray_start=None, ray_dir=None, camera_matrix=None, fov_y=0.0,
                 early_out=1.0 - 1e-5):
        self.W, self.H = W, H
        self.volume = volume
        self.box_min, self.box_size = box_min, box_size
        self.camera_mode = camera_mode
        self.ray_start, self.ray_dir, self.camera_matrix, self.fov_y = ray_start, ray_dir, camera_matrix, fov_y
        self.step_size = step_size
        self.tf_mode, self.tf = tf_mode, tf
        self.blend_mode, self.filter_mode = blend_mode, filter_mode
        self.early_out = early_out

    @property
    def B(self):
        cands = [self.tf.shape[0]]
        if self.filter_mode != capi.FILTER_PRESHADED:
            cands.append(self.volume.shape[0])
        for a in (self.ray_start, self.ray_dir, self.box_min, self.box_size, self.step_size):
            if isinstance(a, np.ndarray):
                cands.append(a.shape[0])
        if isinstance(self.camera_matrix, np.ndarray) and self.camera_matrix.ndim == 3:
            cands.append(self.camera_matrix.shape[0])
        return max(cands)

    def inputs(self, conv=lambda a: a, keep=None) -> capi.Inputs:
        """``conv`` converts each numpy array; converted arrays are appended to ``keep``."""
        keep = keep if keep is not None else []

        def c(a):
            if isinstance(a, np.ndarray):
                r = conv(a)
                keep.append(r)
                return r
            return a

        cm = self.camera_matrix
        if isinstance(cm, np.ndarray) and cm.ndim == 3:
            cm = c(cm)
        elif isinstance(cm, np.ndarray):
            cm = cm.tolist()  # 4x4 scalar variant
        inp = capi.Inputs(
            screen_size=(self.W, self.H), volume=c(self.volume), box_min=c(self.box_min),
            box_size=c(self.box_size), camera_mode=self.camera_mode, step_size=c(self.step_size),
            tf_mode=self.tf_mode, tf=c(self.tf), blend_mode=self.blend_mode,
            volume_filter_mode=self.filter_mode, camera_ray_start=c(self.ray_start),
            camera_ray_dir=c(self.ray_dir), camera_matrix=cm, camera_fov_y_radians=self.fov_y,
            blending_early_out=self.early_out)
        inp._keep = keep
        return inp


def random_tf(tf_mode, R, rng, opacity=8.0):
    C = TF_CHANNELS[tf_mode]
    if tf_mode == capi.TF_IDENTITY:
        return np.array([[[opacity, 0.8]]], dtype=np.float32)
    tf = rng.rand(1, R, C).astype(np.float32)
    tf[..., 3] *= opacity
    if tf_mode == capi.TF_LINEAR:
        pos = np.sort(rng.rand(R)).astype(np.float32)
        pos[0], pos[-1] = 0.0, 1.0
        tf[0, :, 4] = pos
    if tf_mode in (capi.TF_GAUSSIAN, capi.TF_GAUSSIAN_LOG):
        tf[0, :, 4] = np.linspace(0.15, 0.85, R).astype(np.float32)
        tf[0, :, 5] = (0.05 + 0.1 * rng.rand(R)).astype(np.float32)
        if tf_mode == capi.TF_GAUSSIAN_LOG:
            tf[..., :4] = tf[..., :4] * 0.9 + 0.05  # log() needs strictly positive colours
    return tf


def make_scene(*, n=24, W=20, H=16, B=1, camera_mode=capi.CAMERA_REFERENCE_FRAME,
               tf_mode=capi.TF_TEXTURE, R=8, blend_mode=capi.BLEND_BEER_LAMBERT,
               filter_mode=capi.FILTER_TRILINEAR, step_voxels=0.5, seed=0, opacity=8.0,
               volume_kind="ml", x_fastest=False, tensor_scalars=False) -> Scene:
    """Generic small scene: n^3 volume, B fibonacci-sphere views at W x H."""
    rng = np.random.RandomState(seed)
    if volume_kind == "ml":
        vol = synthetic.marschner_lobb(n)
    else:
        vol = rng.rand(1, n, n, n).astype(np.float32)
    if x_fastest:  # the layout of the reference's Volume class (renderer/volume.cpp:21-27)
        vol = np.ascontiguousarray(vol.transpose(0, 3, 2, 1)).transpose(0, 3, 2, 1)
    box_min, box_size = synthetic.default_box(n)
    lat, lon = synthetic.fibonacci_sphere(max(B, 3))
    vp = synthetic.viewport_from_sphere((0, 0, 0), lon[:B], lat[:B], 1.4641, "Ym")
    fov = math.radians(45.0)
    kw = {}
    if camera_mode == capi.CAMERA_REFERENCE_FRAME:
        kw = dict(camera_matrix=vp, fov_y=fov)
    elif camera_mode == capi.CAMERA_RAY_START_DIR:
        rs, rd = synthetic.generate_rays(vp, fov, W, H)
        kw = dict(ray_start=rs, ray_dir=rd)
    else:
        mats = np.stack([synthetic.inverse_view_projection(vp[b, 0], (0, 0, 0), (0, -1, 0), 45.0, W, H)
                         for b in range(B)])
        kw = dict(camera_matrix=mats if (B > 1 or tensor_scalars) else mats[0])
    step = step_voxels / n
    if tensor_scalars:
        box_min = np.array([box_min], dtype=np.float32)
        box_size = np.array([box_size], dtype=np.float32)
        step = np.full((1, 1, 1), step, dtype=np.float32)
    if filter_mode == capi.FILTER_PRESHADED:
        vol = rng.rand(4, n, n, n).astype(np.float32)
        vol[3] *= opacity
        tf_mode = capi.TF_PRESHADED
        tf = np.zeros((1, 1, 1), dtype=np.float32)
    else:
        tf = random_tf(tf_mode, R, rng, opacity)
    return Scene(W=W, H=H, volume=vol, box_min=box_min, box_size=box_size, camera_mode=camera_mode,
                 step_size=step, tf_mode=tf_mode, tf=tf, blend_mode=blend_mode,
                 filter_mode=filter_mode, **kw)


def consistency_tf_scene(tf_mode, R, seed=0, W=128, H=128) -> Scene:
    """unittests/testConsistencyTF.cpp:48-114: 32^3 Marschner-Lobb, inverse view-projection camera
    at (-1,-.5,0) looking at 0, up z, 45 deg, near .1 far 10, step 0.5/64, box [-.5,.5]^3."""
    rng = np.random.RandomState(seed)
    C = TF_CHANNELS[tf_mode]
    tf = rng.rand(1, R, C).astype(np.float32)
    if tf_mode == capi.TF_LINEAR:  # the reference writes i/(R-1) in INTEGER arithmetic (:93-98)
        tf[0, :, 4] = 0.0
        tf[0, R - 1, 4] = 1.0
    cam = synthetic.inverse_view_projection((-1, -0.5, 0), (0, 0, 0), (0, 0, 1), 45.0, W, H, 0.1, 10.0)
    return Scene(W=W, H=H, volume=synthetic.marschner_lobb(32), box_min=(-0.5, -0.5, -0.5),
                 box_size=(1, 1, 1), camera_mode=capi.CAMERA_INVERSE_VIEW_MATRIX, camera_matrix=cam,
                 step_size=0.5 / 64, tf_mode=tf_mode, tf=tf)


def consistency_camera_scene(tf_mode, R, seed=0, W=8, H=8) -> Scene:
    """unittests/testConsistencyCamera.cpp:48-118: same volume, 8x8 per-pixel rays from
    generateRays(viewportFromLookAt((-1,-.5,0), 0, up z), 45 deg), step 0.25."""
    s = consistency_tf_scene(tf_mode, R, seed, W, H)
    vp = synthetic.viewport_from_look_at((-1.0, -0.5, 0.0), (0, 0, 0), (0, 0, 1))
    rs, rd = synthetic.generate_rays(vp, math.radians(45.0), W, H)
    s.camera_mode, s.camera_matrix = capi.CAMERA_RAY_START_DIR, None
    s.ray_start, s.ray_dir = rs, rd
    s.step_size = 0.25
    return s


# ---- running a library on a scene (numpy / host pointers) ---------------------------------------
def np_forward(lib, scene: Scene, dtype=np.float32):
    B, H, W = scene.B, scene.H, scene.W
    inp = scene.inputs(lambda a: a.astype(dtype) if a.dtype != np.int32 else a)
    color = np.zeros((B, H, W, 4), dtype=dtype)
    term = np.zeros((B, H, W), dtype=np.int32)
    lib.render_forward(inp, color, term)
    return color, term


def np_adjoint(lib, scene: Scene, color, term, adj_color, *, volume=False, step=False, camera=False,
               tf=False, dtype=np.float32, step_shape=None, delayed=False):
    B, H, W = scene.B, scene.H, scene.W
    inp = scene.inputs(lambda a: a.astype(dtype) if a.dtype != np.int32 else a)
    out = {}
    if volume:
        out["adj_volume"] = np.zeros(scene.volume.shape, dtype=dtype)
    if step:
        out["adj_step_size"] = np.zeros(step_shape or (B, H, W), dtype=dtype)
    if camera:
        out["adj_camera_ray_start"] = np.zeros((B, H, W, 3), dtype=dtype)
        out["adj_camera_ray_dir"] = np.zeros((B, H, W, 3), dtype=dtype)
    if tf:
        out["adj_tf"] = np.zeros(scene.tf.shape, dtype=dtype)
    adj = capi.make_adjoint_outputs(tf_delayed_accumulation=delayed, **out)
    lib.render_adjoint(inp, color.astype(dtype), term, adj_color.astype(dtype), adj)
    return out


def rel_l2(a, b):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    d = np.linalg.norm((a - b).ravel())
    n = np.linalg.norm(b.ravel())
    return d / n if n > 0 else d


def symmetric_error(a, b):
    """The reference's own consistency metric, unittests/testConsistencyTF.cpp:16-41."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return float(np.max(np.abs(a - b) / np.maximum(1.0, np.abs(a) + np.abs(b)))) if a.size else 0.0


def per_ray_errors(got, want):
    """Per-ray error of a per-ray quantity (leading dims B,H,W), normalised by the RMS per-ray
    magnitude of ``want``. Quantities that contain d(trilinear)/d(position) are DISCONTINUOUS across
    cell faces, so a sample that lands within round-off of a face flips the term for that ray in any
    fp32 implementation (the reference's fp32 and fp64 builds disagree on such rays too); robust
    statistics over rays (median, outlier fraction) separate that from systematic error."""
    got = np.asarray(got, dtype=np.float64)
    want = np.asarray(want, dtype=np.float64)
    n = got.shape[0] * got.shape[1] * got.shape[2]
    g, w = got.reshape(n, -1), want.reshape(n, -1)
    scale = np.sqrt(np.mean(np.sum(w * w, axis=1)))
    return np.sqrt(np.sum((g - w) ** 2, axis=1)) / (scale if scale > 0 else 1.0)


FACE_EPS = 3e-5  # voxels; ~8x the fp32 round-off of a sample coordinate on a <=64^3 test volume


def face_eps(n: int) -> float:
    """8x the fp32 round-off of a sample coordinate in an n^3 volume (n * 2^-24 voxels), never below FACE_EPS."""
    return max(FACE_EPS, 8.0 * n * 2.0 ** -24)


def face_sensitive_rays(scene: Scene, term, eps=None):
    """Boolean (B,H,W) mask of rays with a sample within ``eps`` voxels of a trilinear cell face
    (computed in fp64 by the oracle's diagnostic, oracle/dvr_oracle.cpp face_distance)."""
    import oracle_loader as ol
    if scene.filter_mode != capi.FILTER_TRILINEAR:
        return np.zeros(term.shape, dtype=bool)
    if eps is None:
        eps = face_eps(max(scene.volume.shape[1:]))
    fd = np.ones(term.shape, dtype=np.float64)
    inp = scene.inputs(lambda a: a.astype(np.float64) if a.dtype != np.int32 else a)
    ol.port(64).face_distance(inp, np.ascontiguousarray(term, dtype=np.int32), fd)
    return fd < eps


def assert_position_gradient_close(name, got, want32, want64, tol=1e-4, sensitive=None):
    """Per-ray quantities that contain d(trilinear)/d(position): rel-L2 <= tol against the fp32
    oracle over all rays, or -- because that derivative is DISCONTINUOUS across cell faces, so a
    sample within round-off of a face (always the case for the first sample of a ray that enters
    through a max face of the box) takes either side in any fp32 evaluation order, the reference's
    own fp32 and fp64 builds included --
      * on the rays WITHOUT such a sample (``sensitive`` False): rel-L2 <= tol against the fp64
        oracle, or no further from it than 1.5x the fp32 oracle is;
      * on the face-sensitive rays: median per-ray error no worse than 2x the fp32 oracle's and
        every ray within the size of one flipped sample (half the RMS ray gradient)."""
    import parity_log
    e32 = rel_l2(got, want32)
    if e32 <= tol:
        parity_log.record(name, "strict", rel_l2=e32, tol=tol)
        return
    if callable(want64):  # computed only when the strict gate fails (expensive at the config sizes)
        want64 = want64()
    if callable(sensitive):
        sensitive = sensitive()
    assert sensitive is not None, f"{name}: rel-L2 {e32:.2e} > {tol} and no face mask given"
    parity_log.record(name, "relaxed:face-sensitive-rays", rel_l2=e32, tol=tol,
                      rel_l2_vs_fp64=rel_l2(got, want64), fp32_oracle_noise=rel_l2(want32, want64),
                      masked_fraction=float(np.mean(sensitive)),
                      rel_l2_unmasked_vs_fp64=rel_l2(np.asarray(got)[~sensitive], np.asarray(want64)[~sensitive])
                      if (~sensitive).any() else 0.0)
    lead = sensitive.shape
    g = np.asarray(got, dtype=np.float64).reshape(lead + (-1,))
    w32 = np.asarray(want32, dtype=np.float64).reshape(lead + (-1,))
    w64 = np.asarray(want64, dtype=np.float64).reshape(lead + (-1,))
    ok = ~sensitive
    if ok.any():
        e_got, e_ref = rel_l2(g[ok], w64[ok]), rel_l2(w32[ok], w64[ok])
        assert e_got <= max(tol, 1.5 * e_ref), (
            f"{name}: rel-L2 on face-insensitive rays {e_got:.2e} (fp32 oracle {e_ref:.2e}, {ok.mean():.2f} of rays)")
    if sensitive.any():
        e_got, e_ref = per_ray_errors(got, want64), per_ray_errors(want32, want64)
        m = sensitive.reshape(-1)
        med_got, med_ref = np.median(e_got[m]), np.median(e_ref[m])
        assert med_got <= max(tol / 10, 2 * med_ref), (
            f"{name}: median error on face-sensitive rays {med_got:.2e} (fp32 oracle {med_ref:.2e})")
        assert e_got[m].max() <= 0.5, f"{name}: face-sensitive ray off by {e_got[m].max():.2e} of the RMS gradient"
