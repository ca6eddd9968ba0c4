#!/usr/bin/env python
"""Builds and runs tools/microbench/quad_ceilings.cu on the GPU box and writes the measured gather / RED ceilings
 as JSON: `python tools/microbench/run_ceilings.py [--out gpurun_out/MEASURED_GATHER_RED.json]`.
Copy the file to profiles/MEASURED_GATHER_RED.json (bench.py reads it from there)."""
import argparse
import json
import os
import subprocess
import time

HERE = os.path.dirname(os.path.abspath(__file__))


def clocks():
    try:
        out = subprocess.run(["nvidia-smi", "--query-gpu=name,clocks.sm,clocks.max.sm,clocks_event_reasons.active",
                              "--format=csv,noheader", "-i", "0"], capture_output=True, text=True, timeout=20).stdout
        return out.strip()
    except Exception as e:  # noqa: BLE001
        return f"unavailable: {e}"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(HERE, "..", "..", "gpurun_out", "MEASURED_GATHER_RED.json"))
    ap.add_argument("--layout", default="yzx", choices=["xyz", "yzx"],
                    help="QUAD layout whose ceilings go into by_volume (the other one is kept under other_layout)")
    a = ap.parse_args()
    exe = os.path.join(HERE, "quad_ceilings")
    src = exe + ".cu"
    if not os.path.exists(exe) or os.path.getmtime(exe) < os.path.getmtime(src):
        subprocess.run(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-o", exe, src], check=True)
    p = subprocess.Popen([exe, "both"], stdout=subprocess.PIPE, text=True)
    time.sleep(1.0)
    under_load = clocks()
    out, _ = p.communicate(timeout=600)
    rows = [json.loads(l) for l in out.splitlines() if l.startswith("{")]
    res = {"how": "tools/microbench/quad_ceilings.cu: 2 x LDG.128 / 2 x red.v4 per cell on the QUAD layout, 8x4-ray warp "
                  "tiles, one new cell per iteration, best of 3, harmonic mean over three ray directions",
           "when": time.strftime("%Y-%m-%dT%H:%M:%SZ", time.gmtime()), "clocks_under_load": under_load,
           "layout": a.layout,
           "by_volume": {str(r["n"]): r for r in rows if r.get("layout") == a.layout},
           "other_layout": {str(r["n"]): r for r in rows if r.get("layout") != a.layout}}
    os.makedirs(os.path.dirname(os.path.abspath(a.out)), exist_ok=True)
    with open(a.out, "w") as f:
        json.dump(res, f, indent=1)
    print(json.dumps(res))
    return p.returncode


if __name__ == "__main__":
    raise SystemExit(main())
