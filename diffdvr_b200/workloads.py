"""The five BASELINE.json configurations as synthetic workloads (SURVEY.md 8(d) "C1".."C5").

One definition shared by ``bench.py`` (every config is a bench line), ``tools/bench_configs.py``
(kernel-level timings) and ``tests/test_gpu_config_parity.py`` (parity of the CUDA path against the
reference build at the config sizes).  Producers of INPUTS only; nothing here is on the hot path.

A ``Workload`` keeps its inputs as numpy arrays / python scalars; ``inputs(conv, views)`` makes a
``capi.Inputs`` for any consumer through a conversion callback (identity for the CPU oracle,
``torch.from_numpy(..).cuda()`` for the product).
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Callable, Dict, Optional, Sequence

import numpy as np

from . import capi, synthetic

FOV_Y = math.radians(45.0)
CAMERA_DISTANCE = 1.4641  # 1.1^4, config-files/skull1.json


@dataclass
class Workload:
    name: str
    description: str
    vol_kind: str            # "ml" | "skull"
    n: int                   # volume resolution
    res: int                 # image resolution
    views: int               # views of the whole job
    batch: int               # views per kernel launch
    tf_kind: str             # "identity" | "linear" | "texture" | "gaussian"
    step_voxels: float
    camera: str              # "frame" (ReferenceFrame) | "rays" (RayStartDir from generate_rays)
    adjoint: Sequence[str] = ()        # subset of {"volume", "tf", "stepsize", "camera"}
    forward_mode_D: int = 0            # > 0: forward-mode camera derivatives (C3)
    delayed_tf: bool = False
    _cache: Dict[str, object] = field(default_factory=dict, repr=False)

    # ---- inputs (numpy) ---------------------------------------------------------------------------
    def volume(self) -> np.ndarray:
        if "volume" not in self._cache:
            if self.vol_kind == "ml":
                v = synthetic.marschner_lobb(self.n)
            else:
                v = synthetic.skull(self.n).numpy()
            self._cache["volume"] = np.ascontiguousarray(v, dtype=np.float32)
        return self._cache["volume"]

    def tf(self) -> np.ndarray:
        lin = synthetic.tf_linear_two_peaks()
        return {"identity": synthetic.tf_identity(), "linear": lin,
                "texture": synthetic.tf_texture_from_linear(lin, 64),
                "gaussian": synthetic.tf_gaussian(4)}[self.tf_kind]

    @property
    def tf_mode(self) -> int:
        return {"identity": capi.TF_IDENTITY, "linear": capi.TF_LINEAR, "texture": capi.TF_TEXTURE,
                "gaussian": capi.TF_GAUSSIAN}[self.tf_kind]

    def viewports(self) -> np.ndarray:
        """(views, 3, 3) eye/right/up rows on a Fibonacci sphere (pytests/diffdvr/utils.py:76-89)."""
        lat, lon = synthetic.fibonacci_sphere(max(self.views, 3))
        return synthetic.viewport_from_sphere((0, 0, 0), lon[:self.views], lat[:self.views], CAMERA_DISTANCE, "Ym")

    @property
    def step_size(self) -> float:
        return self.step_voxels / self.n

    def inputs(self, conv: Callable = lambda a: a, views: Optional[Sequence[int]] = None, volume=None, tf=None,
               res: Optional[int] = None) -> capi.Inputs:
        """``views``: indices into viewports() rendered by this call (default: the first launch)."""
        res = res or self.res
        idx = list(views) if views is not None else list(range(min(self.batch, self.views)))
        vp = np.ascontiguousarray(self.viewports()[idx])
        box_min, box_size = synthetic.default_box(self.n)
        kw = dict(screen_size=(res, res), volume=volume if volume is not None else conv(self.volume()),
                  box_min=box_min, box_size=box_size, step_size=self.step_size, tf_mode=self.tf_mode,
                  tf=tf if tf is not None else conv(self.tf()), blend_mode=capi.BLEND_BEER_LAMBERT)
        if self.camera == "frame":
            inp = capi.Inputs(camera_mode=capi.CAMERA_REFERENCE_FRAME, camera_matrix=conv(vp),
                              camera_fov_y_radians=FOV_Y, **kw)
        else:
            rs, rd = synthetic.generate_rays(vp, FOV_Y, res, res)
            inp = capi.Inputs(camera_mode=capi.CAMERA_RAY_START_DIR, camera_ray_start=conv(rs),
                              camera_ray_dir=conv(rd), **kw)
        return inp

    def inputs_from(self, volume, tf, cams, res: Optional[int] = None) -> capi.Inputs:
        """capi.Inputs over tensors the caller already holds: ``cams`` = (viewports,) for the ReferenceFrame camera or
        (ray_start, ray_dir) for per-pixel rays. No host work (the per-call path of bench.py)."""
        res = res or self.res
        box_min, box_size = synthetic.default_box(self.n)
        kw = dict(screen_size=(res, res), volume=volume, box_min=box_min, box_size=box_size, step_size=self.step_size,
                  tf_mode=self.tf_mode, tf=tf, blend_mode=capi.BLEND_BEER_LAMBERT)
        if self.camera == "frame":
            return capi.Inputs(camera_mode=capi.CAMERA_REFERENCE_FRAME, camera_matrix=cams[0],
                               camera_fov_y_radians=FOV_Y, **kw)
        return capi.Inputs(camera_mode=capi.CAMERA_RAY_START_DIR, camera_ray_start=cams[0], camera_ray_dir=cams[1], **kw)

    def summary(self) -> dict:
        return {"workload": f"{self.name}: {self.description}", "views": self.views, "image": [self.res, self.res],
                "volume": [self.n] * 3, "views_per_launch": self.batch, "tf": self.tf_kind,
                "step_voxels": self.step_voxels, "camera": self.camera, "adjoint": list(self.adjoint),
                "forward_mode_D": self.forward_mode_D}


def get(name: str, **override) -> Workload:
    """C1..C5 of BASELINE.json ``configs`` / SURVEY.md 8(d). ``override`` replaces fields (views, res, n, batch)."""
    table = {
        "C1": dict(description="forward render of ML(128) at 256^2, trilinear, piecewise-linear TF, BeerLambert, "
                               "ReferenceFrame camera, step 0.5 voxel",
                   vol_kind="ml", n=128, res=256, views=1, batch=1, tf_kind="linear", step_voxels=0.5, camera="frame"),
        "C2": dict(description="TF reconstruction: SKULL(256), 8 views @256^2 in one launch, Texture TF R=64, "
                               "step 0.2 voxel, forward + L1 loss + adjoint(hasTF, delayed flag)",
                   vol_kind="skull", n=256, res=256, views=8, batch=8, tf_kind="texture", step_voxels=0.2,
                   camera="frame", adjoint=("tf",), delayed_tf=True),
        "C3": dict(description="best viewpoint: SKULL(256), 8 views @512^2, per-pixel rays, Gaussian TF R=4, step 0.2 "
                               "voxel, forward-mode D=6 (ray start / direction) + forward_variables_to_gradients",
                   vol_kind="skull", n=256, res=512, views=8, batch=8, tf_kind="gaussian", step_voxels=0.2,
                   camera="rays", adjoint=("camera",), forward_mode_D=6),
        "C4": dict(description="absorption-only volume reconstruction: SKULL(256), 64 views @512^2, Identity TF [1,1], "
                               "BeerLambert, step 0.2 voxel, forward + L1 loss on alpha + adjoint(hasVolume), "
                               "gradient all-reduce",
                   vol_kind="skull", n=256, res=512, views=64, batch=8, tf_kind="identity", step_voxels=0.2,
                   camera="frame", adjoint=("volume",)),
        "C5": dict(description="large adjoint sweep: SKULL(512), 128 views @1024^2, Texture TF R=64, step 0.5 voxel, "
                               "forward + L1 loss + adjoint(hasVolume + hasTF + hasStepsize), gradient all-reduce",
                   vol_kind="skull", n=512, res=1024, views=128, batch=4, tf_kind="texture", step_voxels=0.5,
                   camera="frame", adjoint=("volume", "tf", "stepsize")),
    }
    if name not in table:
        raise KeyError(f"unknown workload {name!r} (C1..C5)")
    kw = dict(table[name])
    kw.update({k: v for k, v in override.items() if v is not None})
    return Workload(name=name, **kw)


NAMES = ("C1", "C2", "C3", "C4", "C5")
