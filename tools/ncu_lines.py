#!/usr/bin/env python
"""Per-source-line instruction counts of ONE kernel of an ncu report (--import-source on, -lineinfo build):

    python tools/ncu_lines.py report.ncu-rep --launch-skip 9 [--top 40] [--samples N]

Prints the lines of the repository's sources ranked by executed warp instructions (inlined code is attributed to the
line it was written on) and the opcode mix, optionally per marching sample."""
import argparse
import collections
import csv
import io
import re
import subprocess


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("report")
    ap.add_argument("--launch-skip", type=int, default=0)
    ap.add_argument("--top", type=int, default=40)
    ap.add_argument("--samples", type=float, default=0, help="marching samples of the launch (per-sample figures)")
    a = ap.parse_args()
    def page(view):
        raw = subprocess.run(["ncu", "-i", a.report, "--page", "source", "--csv", "--print-source", view,
                              "--launch-skip", str(a.launch_skip), "--launch-count", "1"], capture_output=True,
                             text=True).stdout
        return list(csv.reader(io.StringIO(raw)))

    # 1. the SASS view: every instruction once -> totals and opcode mix
    sass = page("sass")
    print(sass[0][1] if sass and len(sass[0]) > 1 else "?")
    by_op, total, thr = collections.Counter(), 0.0, 0.0
    ie = te = None
    for r in sass:
        if r and r[0] == "Address":
            ie, te = r.index("Instructions Executed"), r.index("Thread Instructions Executed")
            continue
        if ie is None or len(r) <= te:
            continue
        try:
            n, t = float(r[ie]), float(r[te])
        except ValueError:
            continue
        op = r[1].strip().split()
        if op and op[0].startswith("@"):
            op = op[1:]
        by_op[op[0].split(".")[0] if op else "?"] += n
        total += n
        thr += t
    # 2. the CUDA + SASS view: per source line (an inlined instruction is listed under every level of its inline
    #    stack, so the shares of nested helpers overlap with their callers')
    rows = page("cuda,sass")
    # layout: "File Path" / "Function Name" rows, a header row ("Line No", "Source", "Address", "Source", metrics...),
    # then per source line one row with the line's totals followed by its SASS rows (empty "Line No")
    by_line = collections.Counter()
    cur_file, ie, te = None, None, None
    for r in rows:
        if not r:
            continue
        if r[0] == "File Path":
            cur_file = r[1] if len(r) > 1 else None
            continue
        if r[0] == "Line No":
            ie, te = r.index("Instructions Executed"), r.index("Thread Instructions Executed")
            continue
        if ie is None or len(r) <= te:
            continue
        try:
            n, t = float(r[ie]), float(r[te])
        except ValueError:
            continue
        if r[0]:   # a source line: totals of its instructions
            by_line[(cur_file.split("/")[-1] if cur_file else "?", r[0], r[1].strip()[:120])] += n
    print(f"total warp instructions {total:.4g}, thread instructions {thr:.4g}" +
          (f", thread-inst per sample {thr / a.samples:.1f}" if a.samples else ""))
    print("-- opcode mix (warp instructions, share)")
    for op, n in by_op.most_common(25):
        print(f"  {op:10s} {n:12.4g} {100 * n / total:5.1f}%")
    if by_line:
        print("-- source lines by warp instructions")
        for (f, ln, src), n in by_line.most_common(a.top):
            print(f"  {100 * n / total:5.1f}%  {f}:{ln}  {src}")


if __name__ == "__main__":
    main()
