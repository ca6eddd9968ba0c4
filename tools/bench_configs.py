#!/usr/bin/env python
"""Kernel-level throughput of the BASELINE.json configs other than the headline one (bench.py = C4):
Gsamples/s of the forward, adjoint and forward-mode kernels on ONE GPU, CUDA-event timed, L2
flushed between repetitions. Prints one JSON object per config; `--out` also writes a markdown table.

    python tools/bench_configs.py [--configs C1,C2,C3,C5] [--reps 5] [--out profiles/rNN_configs.md]
"""
from __future__ import annotations

import argparse
import json
import math
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402

from diffdvr_b200 import capi, synthetic  # noqa: E402


def timed(fn, reps, flush):
    for _ in range(2):
        fn()
    ms = []
    for _ in range(reps):
        flush.fill_(1.0)
        torch.cuda.synchronize()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        ms.append(s.elapsed_time(e))
    return float(np.median(ms))


def run(name, reps, flush):
    dev = torch.device("cuda", 0)
    lib = capi.load()
    st = lambda: torch.cuda.current_stream().cuda_stream  # noqa: E731
    fov = math.radians(45.0)
    cfg = {
        "C1": dict(vol="ml", n=128, res=256, B=1, tf="linear", step=0.5, cam="frame"),
        "C2": dict(vol="skull", n=256, res=256, B=4, tf="texture", step=0.2, cam="frame"),
        "C3": dict(vol="skull", n=256, res=512, B=8, tf="gaussian", step=0.2, cam="rays"),
        "C5": dict(vol="skull", n=512, res=1024, B=4, tf="texture", step=0.5, cam="frame"),
    }[name]
    n, res, B = cfg["n"], cfg["res"], cfg["B"]
    vol = (torch.from_numpy(synthetic.marschner_lobb(n)) if cfg["vol"] == "ml" else synthetic.skull(n)).to(dev)
    lin = synthetic.tf_linear_two_peaks()
    tf_np, tf_mode = {"linear": (lin, capi.TF_LINEAR), "texture": (synthetic.tf_texture_from_linear(lin, 64), capi.TF_TEXTURE),
                      "gaussian": (synthetic.tf_gaussian(4), capi.TF_GAUSSIAN)}[cfg["tf"]]
    tf = torch.from_numpy(tf_np).to(dev)
    lat, lon = synthetic.fibonacci_sphere(max(B, 3))
    vp_np = synthetic.viewport_from_sphere((0, 0, 0), lon[:B], lat[:B], 1.4641, "Ym")
    box_min, box_size = synthetic.default_box(n)
    kw = dict(screen_size=(res, res), volume=vol, box_min=box_min, box_size=box_size, step_size=cfg["step"] / n,
              tf_mode=tf_mode, tf=tf, blend_mode=capi.BLEND_BEER_LAMBERT)
    if cfg["cam"] == "frame":
        inp = capi.Inputs(camera_mode=capi.CAMERA_REFERENCE_FRAME, camera_matrix=torch.from_numpy(vp_np).to(dev),
                          camera_fov_y_radians=fov, **kw)
    else:
        rs, rd = synthetic.generate_rays(vp_np, fov, res, res)
        inp = capi.Inputs(camera_mode=capi.CAMERA_RAY_START_DIR, camera_ray_start=torch.from_numpy(rs).to(dev),
                          camera_ray_dir=torch.from_numpy(rd).to(dev), **kw)
    color = torch.empty(B, res, res, 4, device=dev)
    term = torch.empty(B, res, res, dtype=torch.int32, device=dev)
    out = {"config": name, "volume": n, "image": res, "views_per_launch": B, "tf": cfg["tf"], "step_voxels": cfg["step"]}
    out["fwd_ms"] = timed(lambda: lib.render_forward(inp, color, term, st()), reps, flush)
    counts = torch.zeros(2, dtype=torch.int64, device=dev)
    lib.count_cell_visits(inp, term, counts, st())
    torch.cuda.synchronize()
    ns, nc = int(counts[0]), int(counts[1])
    out.update(samples=ns, cell_visits=nc, fwd_gsamples=ns / out["fwd_ms"] * 1e-6)
    adj_color = torch.randn(B, res, res, 4, device=dev, generator=torch.Generator(dev).manual_seed(124))
    if name == "C2":
        adj_tf = torch.zeros_like(tf)
        adj = capi.make_adjoint_outputs(adj_tf=adj_tf, tf_delayed_accumulation=True)
        out["adjoint_flags"] = "tf"
    elif name == "C3":
        a_s, a_d = torch.zeros(B, res, res, 3, device=dev), torch.zeros(B, res, res, 3, device=dev)
        adj = capi.make_adjoint_outputs(adj_camera_ray_start=a_s, adj_camera_ray_dir=a_d)
        out["adjoint_flags"] = "camera"
    elif name == "C5":
        adj_vol, adj_tf, adj_step = torch.zeros_like(vol), torch.zeros_like(tf), torch.zeros(B, res, res, device=dev)
        adj = capi.make_adjoint_outputs(adj_volume=adj_vol, adj_tf=adj_tf, adj_step_size=adj_step)
        out["adjoint_flags"] = "volume+tf+stepsize"
    else:
        adj = None
    if adj is not None:
        out["adj_ms"] = timed(lambda: lib.render_adjoint(inp, color, term, adj_color, adj, st()), reps, flush)
        out["adj_gsamples"] = ns / out["adj_ms"] * 1e-6
    if name == "C3":  # forward-mode camera derivatives, D = 6, + the contraction kernel
        D = 6
        grad = torch.empty(B, res, res, D, 4, device=dev)
        s6 = capi.make_fd_settings(D, d_ray_start=(0, 1, 2), d_ray_dir=(3, 4, 5))
        out["fwdmode_ms"] = timed(lambda: lib.render_forward_gradients(inp, s6, color, term, grad, st()), reps, flush)
        out["fwdmode_gsamples"] = ns / out["fwdmode_ms"] * 1e-6
        out["fwdvars_ms"] = timed(lambda: lib.forward_variables_to_gradients(grad, adj_color, s6, adj, st()), reps, flush)
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--configs", default="C1,C2,C3,C5")
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    assert torch.cuda.is_available(), "needs a CUDA device (no CPU fallback)"
    flush = torch.empty(256 * 1024 * 1024 // 4, device="cuda")
    rows = []
    for name in a.configs.split(","):
        r = run(name, a.reps, flush)
        rows.append(r)
        print(json.dumps(r), flush=True)
        torch.cuda.empty_cache()
    if a.out:
        cols = ["config", "volume", "image", "views_per_launch", "tf", "step_voxels", "samples", "fwd_ms", "fwd_gsamples",
                "adjoint_flags", "adj_ms", "adj_gsamples", "fwdmode_ms", "fwdmode_gsamples", "fwdvars_ms"]
        with open(a.out, "w") as f:
            f.write("| " + " | ".join(cols) + " |\n|" + "---|" * len(cols) + "\n")
            for r in rows:
                f.write("| " + " | ".join(f"{r[c]:.3f}" if isinstance(r.get(c), float) else str(r.get(c, "")) for c in cols) + " |\n")


if __name__ == "__main__":
    main()
