#!/bin/bash
# End-of-round evidence on ONE GPU box (through gpurun): GPU tests, smoke, a bench line per BASELINE config, kernel
# timings, ncu --set full of every marching kernel (one launch each), ncu launch list of the C4 bench. Everything lands
# in gpurun_out/ with the tag given (default r02f); summarise here with tools/ncu_summary.py / tools/parity_report.py.
cd "$(dirname "$0")/.." || exit 1
TAG=${1:-r02f}
COMMIT=${2:-unknown}
mkdir -p gpurun_out
timeout 400 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/${TAG}_pytest_gpu.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/${TAG}_smoke.log
NCU="ncu --set full --clock-control none --import-source on -k regex:k_adjoint|k_forward|k_tf_jacobian"
timeout 300 $NCU -f -o gpurun_out/${TAG}a_configs_full python tools/bench_configs.py --profile --configs C1,C2,C3,C4 > gpurun_out/${TAG}a_configs.jsonl 2> gpurun_out/${TAG}a_ncu.err; echo "ncu a rc=$?"
timeout 300 $NCU -f -o gpurun_out/${TAG}b_configs_full python tools/bench_configs.py --profile --configs C5 > gpurun_out/${TAG}b_configs.jsonl 2> gpurun_out/${TAG}b_ncu.err; echo "ncu b rc=$?"
# summarise on the box (ncu -i is here too) so that the bench lines below carry this capture's traffic / bound
python tools/ncu_summary.py gpurun_out/${TAG}a_configs_full.ncu-rep --tag ${TAG}a --commit "$COMMIT" > /dev/null 2> gpurun_out/${TAG}a_summary.err; echo "summary a rc=$?"
python tools/ncu_summary.py gpurun_out/${TAG}b_configs_full.ncu-rep --tag ${TAG}b --commit "$COMMIT" > /dev/null 2> gpurun_out/${TAG}b_summary.err; echo "summary b rc=$?"
python - <<PY
import json, os
out, md = {}, []
for t in ("${TAG}a", "${TAG}b"):
    p = f"profiles/{t}_ncu_kernels.json"
    if os.path.exists(p):
        out.update(json.load(open(p)))
        lines = open(f"profiles/{t}_ncu_kernels.md").read().splitlines()
        md += lines if not md else [l for l in lines if l.startswith("| C")]
if out:
    for v in out.values():
        v["source"] = "profiles/r02_ncu_kernels.md (ncu --set full, one launch)"
    json.dump(out, open("profiles/r02_ncu_kernels.json", "w"), indent=1)
    open("profiles/r02_ncu_kernels.md", "w").write("\n".join(md) + "\n")
    print("merged", sorted(out))
PY
cp profiles/r02_ncu_kernels.json profiles/r02_ncu_kernels.md gpurun_out/
for c in C4 C1 C2 C3 C5; do
  timeout 400 python bench.py --config $c > gpurun_out/${TAG}_bench_${c}_n1.json 2> gpurun_out/${TAG}_bench_${c}_n1.err; echo "bench $c rc=$?"
  head -c 300 gpurun_out/${TAG}_bench_${c}_n1.json; echo
done
timeout 200 python tools/bench_configs.py --reps 7 --out gpurun_out/${TAG}_configs.md > gpurun_out/${TAG}_configs_timed.jsonl 2>/dev/null; echo "configs rc=$?"
timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches_bench_c4.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/${TAG}_bench_under_ncu.json 2> gpurun_out/${TAG}_bench_under_ncu.err; echo "launch list rc=$?"
ls -la gpurun_out | head -40
