#!/usr/bin/env python
"""Diagnostic (GPU box): how far the fp32 TF gradient of a C2 view is from the fp64 reference build -- for the CUDA
kernels, for the reference's own fp32 build, and run to run -- at growing image sizes and for two adjoint colours."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402
import torch  # noqa: E402

import oracle_loader as ol  # noqa: E402
import scenes  # noqa: E402
from diffdvr_b200 import capi, workloads  # noqa: E402


def main():
    lib = capi.load()
    ref32, ref64 = ol.ref(32), ol.ref(64)
    for res in (64, 128, 256):
        w = workloads.get("C2", res=res)
        for kind in ("randn", "sign"):
            inp32 = w.inputs(views=[3])
            c = np.zeros((1, res, res, 4), np.float32)
            t = np.zeros((1, res, res), np.int32)
            ref32.render_forward(inp32, c, t)
            ac = (np.random.RandomState(124).randn(*c.shape).astype(np.float32) if kind == "randn"
                  else (np.sign(c - 0.25) / c[..., 0].size).astype(np.float32))
            outs = {}
            for name, olib, dt in (("ref32", ref32, np.float32), ("ref32_again", ref32, np.float32), ("ref64", ref64, np.float64)):
                cv = lambda a: a.astype(dt) if a.dtype != np.int32 else a  # noqa: E731
                inp = w.inputs(cv, views=[3])
                atf = np.zeros(inp.tf.shape, dt)
                olib.render_adjoint(inp, c.astype(dt), t, ac.astype(dt), capi.make_adjoint_outputs(adj_tf=atf, tf_delayed_accumulation=True))
                outs[name] = atf
            to_dev = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()  # noqa: E731
            inpg = w.inputs(to_dev, views=[3])
            for name in ("gpu", "gpu_again"):
                atf = torch.zeros_like(inpg.tf)
                lib.render_adjoint(inpg, to_dev(c), to_dev(t), to_dev(ac), capi.make_adjoint_outputs(adj_tf=atf, tf_delayed_accumulation=True))
                torch.cuda.synchronize()
                outs[name] = atf.cpu().numpy()
            r = {"res": res, "adj_color": kind, "samples": int(t.sum()),
                 "gpu_vs_ref32": scenes.rel_l2(outs["gpu"], outs["ref32"]), "gpu_vs_ref64": scenes.rel_l2(outs["gpu"], outs["ref64"]),
                 "ref32_vs_ref64": scenes.rel_l2(outs["ref32"], outs["ref64"]),
                 "ref32_run_to_run": scenes.rel_l2(outs["ref32"], outs["ref32_again"]),
                 "gpu_run_to_run": scenes.rel_l2(outs["gpu"], outs["gpu_again"])}
            rows = np.argsort(-np.abs(outs["ref64"][0]).sum(1))[:3]
            r["top_rows"] = {int(k): {"ref64": outs["ref64"][0, k].tolist(), "ref32": outs["ref32"][0, k].tolist(),
                                      "gpu": outs["gpu"][0, k].tolist()} for k in rows}
            print(json.dumps(r), flush=True)


if __name__ == "__main__":
    main()
