#!/bin/bash
# End-of-round evidence on ONE GPU box (through gpurun): GPU tests, smoke, a bench line per BASELINE config, kernel
# timings, ncu --set full of every marching kernel (one launch each), ncu launch list of the C4 bench. Everything lands
# in gpurun_out/ with the tag given (default r02f); summarise here with tools/ncu_summary.py / tools/parity_report.py.
cd "$(dirname "$0")/.." || exit 1
TAG=${1:-r02f}
COMMIT=${2:-unknown}
# PARTIAL="C2 C5": refresh only these configs (ncu entries merged into the committed profiles/r02_ncu_kernels.json, bench
# lines of these configs); no tests, no launch list
PARTIAL=${PARTIAL:-}
mkdir -p gpurun_out
if [ -z "$PARTIAL" ]; then
timeout 400 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/${TAG}_pytest_gpu.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/${TAG}_smoke.log
fi
A_CFGS=C1,C2,C3,C4; B_CFGS=C5; BENCH_CFGS="C4 C1 C2 C3 C5"
if [ -n "$PARTIAL" ]; then A_CFGS=$(echo $PARTIAL | tr ' ' ','); B_CFGS=; BENCH_CFGS=$PARTIAL; fi
NCU="ncu --set full --clock-control none --import-source on -k regex:k_adjoint|k_forward|k_tf_jacobian"
timeout 300 $NCU -f -o gpurun_out/${TAG}a_configs_full python tools/bench_configs.py --profile --configs $A_CFGS > gpurun_out/${TAG}a_configs.jsonl 2> gpurun_out/${TAG}a_ncu.err; echo "ncu a rc=$?"
[ -n "$B_CFGS" ] && timeout 300 $NCU -f -o gpurun_out/${TAG}b_configs_full python tools/bench_configs.py --profile --configs $B_CFGS > gpurun_out/${TAG}b_configs.jsonl 2> gpurun_out/${TAG}b_ncu.err; echo "ncu b rc=$?"
# summarise on the box (ncu -i is here too) so that the bench lines below carry this capture's traffic / bound
python tools/ncu_summary.py gpurun_out/${TAG}a_configs_full.ncu-rep --tag ${TAG}a --commit "$COMMIT" > /dev/null 2> gpurun_out/${TAG}a_summary.err; echo "summary a rc=$?"
[ -n "$B_CFGS" ] && python tools/ncu_summary.py gpurun_out/${TAG}b_configs_full.ncu-rep --tag ${TAG}b --commit "$COMMIT" > /dev/null 2> gpurun_out/${TAG}b_summary.err; echo "summary b rc=$?"
python - <<PY
import json, os
out, md = {}, []
if "${PARTIAL}" and os.path.exists("profiles/r02_ncu_kernels.json"):   # keep the other configs' entries
    out = json.load(open("profiles/r02_ncu_kernels.json"))
    md = open("profiles/r02_ncu_kernels.md").read().splitlines()
for t in ("${TAG}a", "${TAG}b"):
    p = f"profiles/{t}_ncu_kernels.json"
    if os.path.exists(p):
        out.update(json.load(open(p)))
        lines = open(f"profiles/{t}_ncu_kernels.md").read().splitlines()
        new = [l for l in lines if l.startswith("| C")]
        if not md:
            md = lines
        else:
            key = lambda l: l.split("|")[1].strip() if l.startswith("| C") else None
            keys = {key(l) for l in new}
            md = [l for l in md if key(l) not in keys] + new + ([f"\nRows {sorted(keys)}: re-captured at commit $COMMIT."] if "${PARTIAL}" else [])
if out:
    for v in out.values():
        v["source"] = "profiles/r02_ncu_kernels.md (ncu --set full, one launch)"
    json.dump(out, open("profiles/r02_ncu_kernels.json", "w"), indent=1)
    open("profiles/r02_ncu_kernels.md", "w").write("\n".join(md) + "\n")
    print("merged", sorted(out))
PY
cp profiles/r02_ncu_kernels.json profiles/r02_ncu_kernels.md gpurun_out/
for c in $BENCH_CFGS; do
  timeout 400 python bench.py --config $c > gpurun_out/${TAG}_bench_${c}_n1.json 2> gpurun_out/${TAG}_bench_${c}_n1.err; echo "bench $c rc=$?"
  head -c 300 gpurun_out/${TAG}_bench_${c}_n1.json; echo
done
[ -z "$PARTIAL" ] && timeout 200 python tools/bench_configs.py --reps 7 --out gpurun_out/${TAG}_configs.md > gpurun_out/${TAG}_configs_timed.jsonl 2>/dev/null; echo "configs rc=$?"
[ -z "$PARTIAL" ] && timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches_bench_c4.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/${TAG}_bench_under_ncu.json 2> gpurun_out/${TAG}_bench_under_ncu.err; echo "launch list rc=$?"
ls -la gpurun_out | head -40
