#!/usr/bin/env python
"""Summarises an `ncu --set full` report of tools/bench_configs.py --profile (one launch of every marching kernel of the
BASELINE configs) into the two files the repository tracks:

    python tools/ncu_summary.py gpurun_out/rNN_configs_full.ncu-rep --tag rNN [--configs C1,C2,C3,C4,C5]
      -> profiles/rNN_ncu_kernels.json   (read by bench.py: measured DRAM traffic + which unit bounds each kernel)
      -> profiles/rNN_ncu_kernels.md     (the table for people)

The kernels appear in the report in launch order: per config forward, [adjoint], [forward-mode]; the per-sample figures
use the sample counts tools/bench_configs.py printed for the same launches (rNN_configs.jsonl next to the report).
"""
from __future__ import annotations

import argparse
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

METRICS = {
    "ms": ("gpu__time_duration.sum", 1.0),
    "registers": ("launch__registers_per_thread", 1.0),
    "grid": ("launch__grid_size", 1.0),
    "warps_active_pct": ("sm__warps_active.avg.pct_of_peak_sustained_active", 1.0),
    "issue_active_pct": ("smsp__issue_active.avg.pct_of_peak_sustained_active", 1.0),
    "warp_inst": ("smsp__inst_executed.sum", 1.0),
    "lanes_per_inst": ("smsp__thread_inst_executed_per_inst_executed.ratio", 1.0),
    "pipe_fma_pct": ("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", 1.0),
    "pipe_alu_pct": ("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", 1.0),
    "pipe_lsu_pct": ("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", 1.0),
    "pipe_xu_pct": ("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", 1.0),
    "l1tex_pct": ("l1tex__throughput.avg.pct_of_peak_sustained_elapsed", 1.0),
    "lsu_wavefronts_pct": ("l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", 1.0),
    "l2_pct": ("lts__throughput.avg.pct_of_peak_sustained_elapsed", 1.0),
    "dram_pct": ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", 1.0),
    "dram_read_gb": ("dram__bytes_read.sum", 1.0),
    "dram_write_gb": ("dram__bytes_write.sum", 1.0),
    "smem_atom_inst": ("smsp__inst_executed_op_shared_atom.sum", 1.0),
    "smem_atom_bank_conflicts": ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_atom.sum", 1.0),
    "local_ld_sectors": ("l1tex__t_sectors_pipe_lsu_mem_local_op_ld.sum", 1.0),
    "local_st_sectors": ("l1tex__t_sectors_pipe_lsu_mem_local_op_st.sum", 1.0),
}
UNIT_SCALE = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "ms": 1.0, "us": 1e-3, "s": 1e3, "ns": 1e-6}


def read_report(path):
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    out = []
    for r in rows[2:]:
        k = {"kernel": r[idx["Kernel Name"]].replace("void ", "").split("(")[0].strip()}
        for key, (name, _) in METRICS.items():
            if name not in idx:
                k[key] = None
                continue
            txt = r[idx[name]].replace(",", "")
            try:
                v = float(txt)
            except ValueError:
                k[key] = None
                continue
            u = units[idx[name]]
            if key in ("dram_read_gb", "dram_write_gb"):
                v = v * UNIT_SCALE.get(u, 1.0) / 1e9
            elif key == "ms":
                v = v * UNIT_SCALE.get(u, 1.0)
            k[key] = v
        out.append(k)
    return out


def classify(k):
    """Which unit bounds the kernel, read off the capture: the busiest of the issue slots, the L1TEX (LSU wavefronts),
    L2 and DRAM, each as % of its own peak."""
    cands = {"issue": k["issue_active_pct"] or 0, "l1tex": max(k["l1tex_pct"] or 0, k["lsu_wavefronts_pct"] or 0),
             "l2": k["l2_pct"] or 0, "hbm": k["dram_pct"] or 0}
    name = max(cands, key=cands.get)
    if cands[name] < 60.0:
        name = "latency"   # nothing is busy: too few resident warps / a tail
    return name, cands


def pass_of(kernel):
    if kernel.startswith("k_forward_gradients") or kernel.startswith("k_tf_jacobian"):
        return "forward_mode"
    if kernel.startswith("k_adjoint"):
        return "adjoint"
    return "forward"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("report")
    ap.add_argument("--tag", required=True)
    ap.add_argument("--configs-jsonl", default=None, help="output of the same tools/bench_configs.py run (sample counts)")
    ap.add_argument("--commit", default=None)
    a = ap.parse_args()
    kernels = read_report(a.report)
    cfgs = []
    path = a.configs_jsonl or os.path.join(os.path.dirname(a.report), f"{a.tag}_configs.jsonl")
    if os.path.exists(path):
        cfgs = [json.loads(l) for l in open(path) if l.startswith("{")]
    # launch order of bench_configs.py: forward, adjoint (if any), forward-mode (if any) per config
    order = []
    for c in cfgs:
        order.append((c["config"], "forward", c))
        if "adj_ms" in c:
            order.append((c["config"], "adjoint", c))
        if "fwdmode_ms" in c:
            order.append((c["config"], "forward_mode", c))
    commit = a.commit or subprocess.run(["git", "rev-parse", "--short", "HEAD"], cwd=ROOT, capture_output=True,
                                        text=True).stdout.strip()
    result, lines = {}, []
    lines.append(f"# ncu --set full, one launch per marching kernel of the BASELINE configs ({a.tag}, commit {commit})\n")
    lines.append("`ncu --set full --clock-control none --import-source on -k regex:'k_adjoint|k_forward|k_tf_jacobian' python "
                 "tools/bench_configs.py --profile` on one B200; cold caches, serialised launches (durations are for "
                 "shares, not for the bench).\n")
    lines.append("| config:pass | kernel | ms | regs | warps active % | issue active % | lanes/inst | thread-inst/sample | "
                 "FMA / ALU / LSU / XU pipe % | L1TEX (LSU wavefronts) % | L2 % | DRAM % | DRAM GB/launch | smem atomics "
                 "(inst / bank conflicts) | local ld/st sectors | bound |")
    lines.append("|" + "---|" * 16)
    for i, k in enumerate(kernels):
        cfg, pas, c = order[i] if i < len(order) and pass_of(k["kernel"]) == order[i][1] else ("?", pass_of(k["kernel"]), None)
        bound, cands = classify(k)
        per_sample = (k["warp_inst"] * k["lanes_per_inst"] / c["samples"]) if c else None
        dram = (k["dram_read_gb"] or 0) + (k["dram_write_gb"] or 0)
        key = f"{cfg}:{pas}"
        result[key] = dict(k, bound=bound, busiest=cands, thread_inst_per_sample=per_sample,
                           dram_bytes_per_launch=dram * 1e9, commit=commit,
                           source=f"profiles/{a.tag}_ncu_kernels.md (ncu --set full, one launch)")
        ps = f"{per_sample:.0f}" if per_sample else "-"
        lines.append(f"| {key} | `{k['kernel']}` | {k['ms']:.3f} | {k['registers']:.0f} | {k['warps_active_pct']:.1f} | "
                     f"{k['issue_active_pct']:.1f} | {k['lanes_per_inst']:.1f} | {ps} | ")
        lines[-1] += (f"{k['pipe_fma_pct']:.0f} / {k['pipe_alu_pct']:.0f} / {k['pipe_lsu_pct']:.0f} / {k['pipe_xu_pct']:.0f} | "
                      f"{max(k['l1tex_pct'] or 0, k['lsu_wavefronts_pct'] or 0):.0f} | {k['l2_pct']:.0f} | {k['dram_pct']:.0f} | "
                      f"{dram:.2f} | {k['smem_atom_inst'] or 0:.3g} / {k['smem_atom_bank_conflicts'] or 0:.3g} | "
                      f"{k['local_ld_sectors'] or 0:.3g} / {k['local_st_sectors'] or 0:.3g} | {bound} |")
    os.makedirs(os.path.join(ROOT, "profiles"), exist_ok=True)
    with open(os.path.join(ROOT, "profiles", f"{a.tag}_ncu_kernels.json"), "w") as f:
        json.dump(result, f, indent=1)
    with open(os.path.join(ROOT, "profiles", f"{a.tag}_ncu_kernels.md"), "w") as f:
        f.write("\n".join(lines) + "\n")
    print("\n".join(lines))


if __name__ == "__main__":
    sys.exit(main())
