#!/usr/bin/env python
"""Kernel-level throughput of the BASELINE.json configs: Gsamples/s of the forward, adjoint and forward-mode kernels of
ONE launch of each config on ONE GPU, CUDA-event timed, L2 flushed between repetitions, the QUAD copy of the volume held
in a workspace (built once, outside the timed region; the adjoint's gradient scratch is zeroed and folded per call).
Prints one JSON object per config; `--out` also writes a markdown table. `DVR_LIB_PATH` selects an A/B build.

    python tools/bench_configs.py [--configs C1,C2,C3,C4,C5] [--reps 5] [--out profiles/rNN_configs.md]
    ncu --set full -k regex:'k_adjoint|k_forward|k_fwdgrad' ... python tools/bench_configs.py --profile
"""
from __future__ import annotations

import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402

from diffdvr_b200 import capi, workloads  # noqa: E402


WARMUP = 2


def timed(fn, reps, flush):
    for _ in range(WARMUP):
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


def run(name, reps, flush, adjoint_override=None):
    dev = torch.device("cuda", 0)
    lib = capi.load()
    w = workloads.get(name)
    B, res = min(w.batch, w.views), w.res

    def to_dev(a):
        return torch.from_numpy(np.ascontiguousarray(a)).to(dev)

    inp = w.inputs(to_dev, views=list(range(B)))
    ws = lib.workspace()
    ws.set_volume(inp.volume)
    opts = capi.make_call_options(workspace=ws)
    color = torch.empty(B, res, res, 4, device=dev)
    term = torch.empty(B, res, res, dtype=torch.int32, device=dev)
    out = {"config": name, "volume": w.n, "image": res, "views_per_launch": B, "tf": w.tf_kind, "step_voxels": w.step_voxels,
           "lib": os.path.basename(capi.library_path())}
    out["fwd_ms"] = timed(lambda: lib.render_forward(inp, color, term, options=opts), reps, flush)
    counts = torch.zeros(2, dtype=torch.int64, device=dev)
    lib.count_cell_visits(inp, term, counts)
    torch.cuda.synchronize()
    ns, nc = int(counts[0]), int(counts[1])
    out.update(samples=ns, cell_visits=nc, fwd_gsamples=ns / out["fwd_ms"] * 1e-6)
    adj_color = torch.randn(B, res, res, 4, device=dev, generator=torch.Generator(dev).manual_seed(124))
    flags = adjoint_override if adjoint_override is not None else w.adjoint
    bufs = {}
    if "volume" in flags:
        bufs["adj_volume"] = torch.zeros_like(inp.volume)
    if "tf" in flags:
        bufs["adj_tf"] = torch.zeros_like(inp.tf)
    if "stepsize" in flags:
        bufs["adj_step_size"] = torch.zeros(B, res, res, device=dev)
    if "camera" in flags:
        bufs["adj_camera_ray_start"] = torch.zeros(B, res, res, 3, device=dev)
        bufs["adj_camera_ray_dir"] = torch.zeros(B, res, res, 3, device=dev)
    if bufs:
        adj = capi.make_adjoint_outputs(tf_delayed_accumulation=w.delayed_tf, **bufs)
        out["adjoint_flags"] = "+".join(flags)
        out["adj_ms"] = timed(lambda: lib.render_adjoint(inp, color, term, adj_color, adj, options=opts), reps, flush)
        out["adj_gsamples"] = ns / out["adj_ms"] * 1e-6
    if w.forward_mode_D:  # forward-mode camera derivatives + the contraction kernel
        D = w.forward_mode_D
        grad = torch.empty(B, res, res, D, 4, device=dev)
        s6 = capi.make_fd_settings(D, d_ray_start=(0, 1, 2), d_ray_dir=(3, 4, 5))
        out["fwdmode_ms"] = timed(lambda: lib.render_forward_gradients(inp, s6, color, term, grad, options=opts), reps, flush)
        out["fwdmode_gsamples"] = ns / out["fwdmode_ms"] * 1e-6
        out["fwdvars_ms"] = timed(lambda: lib.forward_variables_to_gradients(grad, adj_color, s6, adj), reps, flush)
    ws.close()
    return out


def run_fwdmode_tf(reps, flush):
    """Forward-mode with TF tangents: one view of the C2 scene (SKULL(256), 256^2, Texture R=64 ->
    D = 256) and a D = 32 table (Texture R=8, the reference's testConsistencyTF.cpp shape), the one-pass sample-major
    kernel against the slot-major ceil(D/8) passes it replaces."""
    dev = torch.device("cuda", 0)
    lib = capi.load()
    rows = []
    for label, R in (("C2 scene, Texture R=64 (D=256)", 64), ("C2 scene, Texture R=8 (D=32)", 8),
                     ("C3 scene (one 512^2 view), Gaussian R=4 (D=24)", -4)):
        if R > 0:
            w = workloads.get("C2", views=1, batch=1)
            tf = torch.from_numpy(np.ascontiguousarray(
                workloads.synthetic.tf_texture_from_linear(workloads.synthetic.tf_linear_two_peaks(), R))).to(dev)
            C = 4
        else:
            w = workloads.get("C3", views=1, batch=1)
            R, C = -R, 6
            tf = torch.from_numpy(np.ascontiguousarray(w.tf())).to(dev)
        inp = w.inputs(lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev), views=[0], tf=tf)
        ws = lib.workspace()
        ws.set_volume(inp.volume)
        D = R * C
        d_tf = torch.arange(D, dtype=torch.int32, device=dev).reshape(1, R, C)
        st = capi.make_fd_settings(D, d_tf=d_tf, has_tf_derivatives=True)
        color = torch.empty(1, w.res, w.res, 4, device=dev)
        term = torch.empty(1, w.res, w.res, dtype=torch.int32, device=dev)
        grad = torch.empty(1, w.res, w.res, D, 4, device=dev)
        out = {"case": label, "D": D, "jacobian_mib": grad.numel() * 4 / 2**20}
        res = {}
        for name, mode in (("one_pass", 0), ("slot_major", 1)):
            opts = capi.make_call_options(workspace=ws, forward_mode_tf=mode)
            n0 = lib.launch_count()
            lib.render_forward_gradients(inp, st, color, term, grad, options=opts)
            torch.cuda.synchronize()
            out[name + "_launches"] = lib.launch_count() - n0
            out[name + "_ms"] = timed(lambda: lib.render_forward_gradients(inp, st, color, term, grad, options=opts),
                                      reps, flush)
            res[name] = grad.clone()
        out["speedup"] = out["slot_major_ms"] / out["one_pass_ms"]
        out["rel_l2_between_them"] = float((res["one_pass"] - res["slot_major"]).norm() / res["slot_major"].norm())
        ws.close()
        del res
        torch.cuda.empty_cache()
        rows.append(out)
        print(json.dumps(out), flush=True)
    return rows


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--configs", default="C1,C2,C3,C4,C5")
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--out", default=None)
    ap.add_argument("--profile", action="store_true",
                    help="launch every kernel exactly once (no warm-up, one repetition): the run to put under ncu")
    ap.add_argument("--fwdmode-tf", action="store_true", help="forward-mode TF tangents: one pass vs slot-major passes")
    a = ap.parse_args()
    if a.profile:
        global WARMUP
        WARMUP, a.reps = 0, 1
    assert torch.cuda.is_available(), "needs a CUDA device (no CPU fallback)"
    flush = torch.empty(256 * 1024 * 1024 // 4, device="cuda")
    if a.fwdmode_tf:
        run_fwdmode_tf(a.reps, flush)
        return
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
