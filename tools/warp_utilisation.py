#!/usr/bin/env python
"""How much of the marchers' lane idleness comes from UNEQUAL TRIP COUNTS inside a warp (the part a persistent
"refill idle lanes with new rays" scheme could recover)? For the bench workload (C4 views, 8 x 4 pixel warp tiles as
in thread_pixel()) prints sum(trip counts) / (32 * max trip count in the warp), weighted by work.

    python tools/warp_utilisation.py > profiles/r01_warp_utilisation.md
"""
import math
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from diffdvr_b200 import capi, synthetic  # noqa: E402


def main():
    sys.argv = [sys.argv[0]]
    a = bench.parse_args()
    lib = capi.load()
    dev = torch.device("cuda", 0)
    vol = (0.5 * synthetic.skull(a.vol) + 0.05).to(dev)
    vps = torch.from_numpy(bench.make_views(a))[: a.batch].contiguous().to(dev)
    box_min, box_size = synthetic.default_box(a.vol)
    tf = torch.from_numpy(synthetic.tf_identity()).to(dev)
    H = W = a.res
    inp = capi.Inputs(screen_size=(W, H), volume=vol, box_min=box_min, box_size=box_size,
                      camera_mode=capi.CAMERA_REFERENCE_FRAME, camera_matrix=vps, camera_fov_y_radians=math.radians(45.0),
                      step_size=a.step_voxels / a.vol, tf_mode=capi.TF_IDENTITY, tf=tf,
                      blend_mode=capi.BLEND_BEER_LAMBERT)
    color = torch.empty(len(vps), H, W, 4, device=dev)
    term = torch.empty(len(vps), H, W, dtype=torch.int32, device=dev)
    lib.render_forward(inp, color, term, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    t = term.cpu().numpy().astype(np.int64)
    B = t.shape[0]
    print(f"| tile (w x h) | warps with work | lane utilisation by trip count | same, CTA-level (128 rays) |")
    print("|---|---|---|---|")
    for tw, th in ((8, 4), (4, 8), (16, 2), (32, 1)):
        tiles = t.reshape(B, H // th, th, W // tw, tw).transpose(0, 1, 3, 2, 4).reshape(-1, tw * th)
        mx = tiles.max(axis=1)
        util = tiles.sum() / (32.0 * mx.sum())
        cta = t.reshape(B, H // 8, 8, W // 16, 16).transpose(0, 1, 3, 2, 4).reshape(-1, 128)
        cta_util = cta.sum() / (128.0 * cta.max(axis=1).sum())
        print(f"| {tw} x {th} | {(mx > 0).mean():.3f} | {util:.3f} | {cta_util:.3f} |")
    print(f"\nsamples per launch: {t.sum()}, rays: {t.size}, rays with work: {(t > 0).mean():.3f}, "
          f"mean trip count of those: {t[t > 0].mean():.1f}, max: {t.max()}")


if __name__ == "__main__":
    main()
