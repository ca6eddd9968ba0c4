#!/usr/bin/env python
"""Benchmark of the hot path on B200: the 256^3 volume-gradient step (BASELINE.json config 4,
SURVEY.md 8(d) "C4"): SKULL(256) synthetic density, 64 views at 512^2 on a Fibonacci sphere,
Identity TF [1,1], Beer-Lambert, step 0.2 voxel; every view = forward render + adjoint pass
scattering d(loss)/d(volume); views are sharded over the GPUs of one box and the volume gradient
is summed with one NCCL all-reduce per step.

    python bench.py --gpus N --steps K --warmup W            # our CUDA path
    python bench.py --impl reference --steps K --warmup W    # the reference's CPU path (oracle)

Prints ONE JSON line (see the keys at the bottom). value = views/s of the whole job.
"""
from __future__ import annotations

import argparse
import json
import math
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

METRIC = "views/s, 256^3 volume-gradient step (forward + adjoint(volume) + gradient all-reduce)"
UNIT = "views/s"


def parse_args():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=3)
    p.add_argument("--warmup", type=int, default=3)
    p.add_argument("--impl", default="ours", choices=["ours", "reference"])
    p.add_argument("--views", type=int, default=64)
    p.add_argument("--res", type=int, default=512)
    p.add_argument("--vol", type=int, default=256)
    p.add_argument("--batch", type=int, default=8, help="views per kernel launch")
    p.add_argument("--step-voxels", type=float, default=0.2)
    p.add_argument("--no-cpu-baseline", action="store_true")
    return p.parse_args()


def workload_config(a):
    return {
        "workload": f"C4: SKULL({a.vol}) {a.views} views @{a.res}^2, Identity TF [1,1], BeerLambert, "
                    f"trilinear, step {a.step_voxels} voxel, ReferenceFrame camera (fibonacci sphere, "
                    f"dist 1.4641, fovY 45deg), L1 loss on alpha, adjoint(hasVolume)",
        "views": a.views, "image": [a.res, a.res], "volume": [a.vol] * 3, "views_per_launch": a.batch,
        "sharding": "views round-robin over ranks, volume replicated, adj_volume all-reduce(sum)",
    }


def make_views(a):
    from diffdvr_b200 import synthetic
    lat, lon = synthetic.fibonacci_sphere(a.views)
    return synthetic.viewport_from_sphere((0, 0, 0), lon, lat, 1.4641, "Ym")  # (V,3,3)


# ------------------------------------------------------------------------------------------------
# clocks sampling (B200_PROFILING.md "clocks DURING the timed region")
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    def __init__(self, index):
        self.samples, self.reasons, self._stop = [], set(), threading.Event()
        self.max_mhz = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None
        self.t = threading.Thread(target=self._run, daemon=True)

    def _run(self):
        nv = self.nv
        names = {
            "hw_slowdown": getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8),
            "hw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40),
            "sw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20),
            "sw_power_cap": getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4),
        }
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    mask = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if mask & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            self._stop.wait(0.05)

    def __enter__(self):
        if self.nv:
            self.t.start()
        return self

    def __exit__(self, *exc):
        self._stop.set()
        if self.nv:
            self.t.join(timeout=2)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["unavailable"]}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons)}


# ------------------------------------------------------------------------------------------------
# CPU arm: the reference's own host kernels (oracle/_ref) or the CPU port, timed on host cores
# ------------------------------------------------------------------------------------------------
def cpu_view_time(a, res, vol_np, ref_alpha_fn=None):
    """Times forward + adjoint(volume) of ONE view at res x res on the CPU oracle; returns
    (seconds, kind, lib)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_loader
    from diffdvr_b200 import capi, synthetic
    lib = oracle_loader.ref(32)
    kind = "reference"
    if lib is None:
        lib, kind = oracle_loader.port(32), "port"
    vp = make_views(a)[:1].copy()
    box_min, box_size = synthetic.default_box(a.vol)
    inp = capi.Inputs(screen_size=(res, res), volume=vol_np, box_min=box_min, box_size=box_size,
                      camera_mode=capi.CAMERA_REFERENCE_FRAME, camera_matrix=vp,
                      camera_fov_y_radians=math.radians(45.0), step_size=a.step_voxels / a.vol,
                      tf_mode=capi.TF_IDENTITY, tf=synthetic.tf_identity(), blend_mode=capi.BLEND_BEER_LAMBERT)
    color = np.zeros((1, res, res, 4), np.float32)
    term = np.zeros((1, res, res), np.int32)
    adj_vol = np.zeros_like(vol_np)
    adj = capi.make_adjoint_outputs(adj_volume=adj_vol)
    t0 = time.perf_counter()
    lib.render_forward(inp, color, term)
    adj_color = np.zeros_like(color)
    adj_color[..., 3] = np.sign(color[..., 3] - 0.25) / color[..., 3].size
    lib.render_adjoint(inp, color, term, adj_color, adj)
    return time.perf_counter() - t0, kind, int(term.sum())


def cpu_baseline(a, vol_np, budget_s=20.0):
    t_small, kind, _ = cpu_view_time(a, 64, vol_np)
    res = 64
    for cand in (128, 256, 512):
        if cand <= a.res and t_small * (cand / 64.0) ** 2 <= budget_s:
            res = cand
    t, kind, nsamp = (t_small, kind, 0) if res == 64 else cpu_view_time(a, res, vol_np)
    views_per_s = (res * res) / float(a.res * a.res) / t
    return {"value": views_per_s, "unit": UNIT, "cores": os.cpu_count(), "kind": kind,
            "sample": f"1 view forward+adjoint(volume) at {res}x{res} of the same {a.vol}^3 volume/step, "
                      f"{t:.2f} s, scaled by ({res}/{a.res})^2 rays to one {a.res}^2 view",
            "seconds": t}


def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from diffdvr_b200 import synthetic
    vol_np = (0.5 * synthetic.skull(a.vol).numpy() + 0.05).astype(np.float32)
    # bounded sample per step: one view at a resolution that keeps a step to a few seconds
    t_probe, kind, _ = cpu_view_time(a, 64, vol_np)
    res = 64
    for cand in (128, 256):
        if t_probe * (cand / 64.0) ** 2 <= 8.0:
            res = cand
    for _ in range(a.warmup):
        cpu_view_time(a, res, vol_np)
    times = [cpu_view_time(a, res, vol_np)[0] for _ in range(a.steps)]
    t = float(np.mean(times))
    value = (res * res) / float(a.res * a.res) / t
    sample = (f"each step = 1 view forward+adjoint(volume) at {res}x{res} (same volume/step size), "
              f"scaled by ({res}/{a.res})^2 rays to {a.res}^2 views")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps,
        "warmup": a.warmup, "ms_per_step": 1e3 * t, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": workload_config(a),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": os.cpu_count(), "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


# ------------------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------------------
def run_ours(a):
    import torch
    import torch.distributed as dist
    from diffdvr_b200 import capi, multigpu, synthetic

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device: the product has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    lib = capi.load()
    stream = lambda: torch.cuda.current_stream().cuda_stream  # noqa: E731

    # ---- synthetic inputs -----------------------------------------------------------------------
    truth = synthetic.skull(a.vol)                      # (1,n,n,n) target volume
    vol_host = (0.5 * truth + 0.05).contiguous().pin_memory()   # the volume being optimised
    viewports = torch.from_numpy(make_views(a))
    my_views = multigpu.shard_views(a.views, rank, world)
    launches = list(multigpu.batches(my_views, a.batch))
    vp_host = [viewports[idx].contiguous().pin_memory() for idx in launches]
    box_min, box_size = synthetic.default_box(a.vol)
    tf = torch.from_numpy(synthetic.tf_identity()).to(dev)
    fov, step = math.radians(45.0), a.step_voxels / a.vol
    H = W = a.res

    def inputs(volume, vp):
        return capi.Inputs(screen_size=(W, H), volume=volume, box_min=box_min, box_size=box_size,
                           camera_mode=capi.CAMERA_REFERENCE_FRAME, camera_matrix=vp, camera_fov_y_radians=fov,
                           step_size=step, tf_mode=capi.TF_IDENTITY, tf=tf, blend_mode=capi.BLEND_BEER_LAMBERT)

    # target alpha images from the ground-truth volume (setup, untimed)
    truth_dev = truth.to(dev)
    target_alpha = []
    for vp in vp_host:
        v = vp.to(dev)
        color = torch.empty(len(v), H, W, 4, device=dev)
        term = torch.empty(len(v), H, W, dtype=torch.int32, device=dev)
        lib.render_forward(inputs(truth_dev, v), color, term, stream())
        target_alpha.append(color[..., 3].clone())
    del truth_dev

    volume = vol_host.to(dev)
    vp_dev = [v.to(dev) for v in vp_host]
    adj_volume = torch.zeros_like(volume)
    grad_host = torch.empty_like(vol_host).pin_memory()
    colors = [torch.empty(len(v), H, W, 4, device=dev) for v in vp_dev]
    terms = [torch.empty(len(v), H, W, dtype=torch.int32, device=dev) for v in vp_dev]
    n_pix_total = float(a.views * H * W)
    fwd_ev, adj_ev = [], []

    def step_fn(e2e: bool, record: bool):
        """One volume-gradient step over this rank's views. Returns the loss tensor."""
        vol = volume
        vps = vp_dev
        if e2e:  # host -> device copies of this step's inputs (pinned memory)
            vol = vol_host.to(dev, non_blocking=True)
            vps = [v.to(dev, non_blocking=True) for v in vp_host]
        adj_volume.zero_()
        loss = torch.zeros((), device=dev)
        for k, vp in enumerate(vps):
            inp = inputs(vol, vp)
            if record:
                e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
                e0.record()
            lib.render_forward(inp, colors[k], terms[k], stream())
            if record:
                e1.record()
            diff = colors[k][..., 3] - target_alpha[k]
            loss += diff.abs().sum() / n_pix_total          # L1 on alpha (compare_reconstruction.py)
            adj_color = torch.zeros_like(colors[k])
            adj_color[..., 3] = torch.sign(diff) / n_pix_total
            adj = capi.make_adjoint_outputs(adj_volume=adj_volume)
            if record:
                e1b = torch.cuda.Event(enable_timing=True)
                e1b.record()
            lib.render_adjoint(inp, colors[k], terms[k], adj_color, adj, stream())
            if record:
                e2.record()
                fwd_ev.append((e0, e1))
                adj_ev.append((e1b, e2))
        multigpu.allreduce_gradients([adj_volume])        # no-op on one GPU
        if world > 1:
            dist.all_reduce(loss)
        if e2e:  # device -> host read of the step's result
            grad_host.copy_(adj_volume, non_blocking=True)
            return float(loss.item())
        return loss

    flush_buf = torch.empty(256 * 1024 * 1024 // 4, device=dev)

    def timed(e2e: bool, steps: int, record: bool):
        """Per-step CUDA-event timing with an L2 flush (256 MiB write) between timed steps."""
        total_ms = 0.0
        for _ in range(steps):
            flush_buf.fill_(1.0)
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            step_fn(e2e, record)
            e.record()
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()
            ms = torch.tensor([s.elapsed_time(e)], device=dev)
            if world > 1:
                dist.all_reduce(ms, op=dist.ReduceOp.MAX)   # max over ranks
            total_ms += float(ms.item())
        return total_ms / steps

    # ---- warm-up, then the timed regions -----------------------------------------------------------
    for _ in range(max(a.warmup, 3)):
        step_fn(False, False)
    torch.cuda.synchronize()
    n0 = lib.launch_count()
    with ClockSampler(local_rank) as clocks:
        ms_step = timed(False, a.steps, True)
        launches_timed = lib.launch_count() - n0
        ms_e2e = timed(True, a.steps, False)
    torch.cuda.synchronize()

    # ---- accounting: samples, cell visits, algorithmic bytes -----------------------
    counts = torch.zeros(2, dtype=torch.int64, device=dev)
    for k, vp in enumerate(vp_dev):
        lib.count_cell_visits(inputs(volume, vp), terms[k], counts, stream())
    torch.cuda.synchronize()
    if world > 1:
        dist.all_reduce(counts)
    n_samples, n_cells = int(counts[0]), int(counts[1])
    fwd_ms = float(np.mean([s.elapsed_time(e) for s, e in fwd_ev]))
    adj_ms = float(np.mean([s.elapsed_time(e) for s, e in adj_ev]))
    n_launch_total = a.views / float(a.batch)              # adjoint launches per step, all ranks
    rays_per_launch = a.batch * H * W
    cells_per_launch = n_cells / n_launch_total
    samples_per_launch = n_samples / n_launch_total
    # adjoint: 32 B corner reads + 64 B read-modify-write of 8 gradient corners per cell visit,
    # + per ray: colour 16 + terminationIndex 4 + adj_color 16
    adj_bytes = 96.0 * cells_per_launch + rays_per_launch * 36.0
    fwd_bytes = 32.0 * cells_per_launch + rays_per_launch * 20.0
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    traffic = None
    try:
        traffic = json.load(open(os.path.join(ROOT, "profiles", "adjoint_traffic.json"))).get("dram_bytes_per_launch")
    except Exception:
        pass
    achieved = adj_bytes / (adj_ms * 1e-3) / 1e9

    if rank == 0:
        out = {
            "metric": METRIC, "value": a.views / (ms_step * 1e-3), "unit": UNIT, "n_gpus": world,
            "steps": a.steps, "warmup": max(a.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": dict(workload_config(a), l2="256 MiB L2 flush between timed steps; per-step working set "
                           "(volume 64 MiB + its QUAD copy 256 MiB + QUAD gradient scratch 256 MiB + adj_volume 64 MiB + images) "
                           "also exceeds the 126 MB L2"),
            "fwd_gsamples_per_s": samples_per_launch / (fwd_ms * 1e-3) / 1e9,
            "adj_gsamples_per_s": samples_per_launch / (adj_ms * 1e-3) / 1e9,
            "fwd_ms_per_launch": fwd_ms, "adj_ms_per_launch": adj_ms,
            "samples_per_step": n_samples, "cell_visits_per_step": n_cells,
            "roofline": {"kernel": "k_adjoint<TrilinearQuad,Identity,BeerLambert,hasVol>", "bound": "hbm",
                         "duration": "CUDA events around the whole dvr_render_adjoint call: QUAD copy + memset + "
                                     "march + gradient fold (the march is 97 % of it, profiles/r01_summary.md)",
                         "achieved": achieved, "peak": peak,
                         "peak_source": "MEASURED_PEAKS.json hbm_gbs (measured)" if peaks else "fallback 6650",
                         "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                         "algorithmic_bytes_per_launch": adj_bytes,
                         "forward": {"achieved": fwd_bytes / (fwd_ms * 1e-3) / 1e9,
                                     "frac": fwd_bytes / (fwd_ms * 1e-3) / 1e9 / peak,
                                     "algorithmic_bytes_per_launch": fwd_bytes}},
            "e2e": {"value": a.views / (ms_e2e * 1e-3), "unit": UNIT, "ms_per_step": ms_e2e,
                    "h2d_bytes_per_step": int(vol_host.numel() * 4 + sum(v.numel() * 4 for v in vp_host)),
                    "d2h_bytes_per_step": int(grad_host.numel() * 4 + 4)},
            "gpu_launches": int(launches_timed),
            "clocks": clocks.summary(),
        }
        if world == 1 and not a.no_cpu_baseline:
            out["cpu_baseline"] = cpu_baseline(a, vol_host.numpy())
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


def main():
    a = parse_args()
    # the contract is ONE JSON line on stdout: libraries (NCCL prints its version banner there) get
    # stderr as their stdout for the whole run, the JSON line goes to the real stdout at the end
    sys.stdout.flush()
    real_stdout = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    sys.stdout = real_stdout
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
    real_stdout.flush()


if __name__ == "__main__":
    main()
