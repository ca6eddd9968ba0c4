#!/usr/bin/env python
"""Kernel-level timing of the callers either side of the marcher at C3 size
(8 views @512^2, D = 6; 256^3 volume): Camera::generateRays forward / backward, Renderer::compareToImage and
renderer::Volume's producers (implicit data sets, .cvol body conversion, mipmap pooling).
All three are streaming kernels; the figure of merit is algorithmic bytes / time against the measured
HBM copy bandwidth (MEASURED_PEAKS.json hbm_gbs). 8 x 512^2 x 24 B = 50 MB per ray tensor fits the
126 MB L2, so a 256 MiB buffer is written between timed launches to flush it.

    python tools/bench_aux.py > profiles/r01_aux_kernels.md
"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import camera_cases as cc  # noqa: E402
from diffdvr_b200 import capi  # noqa: E402


def timed(fn, flush, reps=10):
    ts = []
    for _ in range(3):
        fn()
    for _ in range(reps):
        flush.fill_(1.0)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        e1.synchronize()
        ts.append(e0.elapsed_time(e1))
    return float(np.median(ts))


def main():
    lib = capi.load()
    peak = 6500.0
    try:
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        pass
    B, H, W, D, fov = 8, 512, 512, 6, np.radians(45.0)
    flush = torch.empty(64 << 20, device="cuda")
    vp = torch.from_numpy(cc.make_viewports(B, np.random.RandomState(5))).cuda()
    start, direction = torch.empty(B, H, W, 3, device="cuda"), torch.empty(B, H, W, 3, device="cuda")
    gs, gd = torch.randn(B, H, W, 3, device="cuda"), torch.randn(B, H, W, 3, device="cuda")
    adj = torch.empty(B, 3, 3, device="cuda")
    color, ref = torch.rand(B, H, W, 4, device="cuda"), torch.rand(B, H, W, 4, device="cuda")
    grads = torch.randn(B, H, W, D, 4, device="cuda")
    diff, gout = torch.empty(B, H, W, 1, device="cuda"), torch.empty(B, H, W, D, device="cuda")
    dr, gr = torch.empty(1, device="cuda"), torch.empty(D, device="cuda")
    s = torch.cuda.current_stream().cuda_stream
    pix = B * H * W
    rows = [
        ("generate_rays (forward)", lambda: lib.generate_rays(vp, fov, start, direction, s), pix * 24),
        ("generate_rays (backward, 2 launches)", lambda: lib.generate_rays_adjoint(vp, fov, gs, gd, adj, s), pix * 24),
        (f"compare_to_image per pixel, D={D}", lambda: lib.compare_to_image(color, grads, ref, diff, gout, False, s),
         pix * (32 + 16 * D + 4 + 4 * D)),
        (f"compare_to_image reduced, D={D} (2 launches)", lambda: lib.compare_to_image(color, grads, ref, dr, gr, True, s),
         pix * (32 + 16 * D)),
    ]
    # renderer::Volume producers (row f3) at the configs' volume size
    n = 256
    vox = n ** 3
    vol, vol2 = torch.empty(1, n, n, n, device="cuda"), torch.empty(1, n, n, n, device="cuda")
    raw8 = torch.randint(0, 256, (vox,), dtype=torch.uint8, device="cuda")
    raw32 = torch.rand(vox, device="cuda")
    half = torch.empty(1, n // 2, n // 2, n // 2, device="cuda")
    rows += [
        (f"create_implicit_volume Sphere {n}^3", lambda: lib.create_implicit_volume(2, vol, None, s), vox * 4),
        (f"create_implicit_volume MarschnerLobb {n}^3 (fp64 sin/cos)", lambda: lib.create_implicit_volume(0, vol, None, s), vox * 4),
        (f"volume_from_raw u8 {n}^3 (transpose + convert)", lambda: lib.volume_from_raw(raw8, 0, vol, s), vox * 5),
        (f"volume_from_raw f32 {n}^3 (transpose)", lambda: lib.volume_from_raw(raw32, 2, vol, s), vox * 8),
        (f"volume_to_raw {n}^3 (transpose)", lambda: lib.volume_to_raw(vol, raw32, s), vox * 8),
        (f"downsample_volume {n}^3 -> {n // 2}^3", lambda: lib.downsample_volume(vol, half, s), vox * 4 + vox // 2),
    ]
    print(f"| kernel ({B} views @{W}x{H}, L2 flushed between launches) | ms | algorithmic MB | GB/s | of {peak:.0f} GB/s HBM copy peak |")
    print("|---|---|---|---|---|")
    for name, fn, nbytes in rows:
        ms = timed(fn, flush)
        gbs = nbytes / ms * 1e-6
        print(f"| {name} | {ms:.4f} | {nbytes / 1e6:.1f} | {gbs:.0f} | {gbs / peak:.2f} |")


if __name__ == "__main__":
    main()
