#!/bin/bash
# A/B of library builds on one GPU box: tools/ab_run.sh <suffix> [<suffix> ...] runs tools/bench_configs.py on
# diffdvr_b200/libdvr_b200_<suffix>.so (csrc/Makefile OBJ=... LIB=... EXTRA=...) and writes gpurun_out/ab_<suffix>.{md,jsonl}
cd "$(dirname "$0")/.." || exit 1
mkdir -p gpurun_out
for v in "$@"; do
  lib=$PWD/diffdvr_b200/libdvr_b200_$v.so
  [ "$v" = default ] && lib=$PWD/diffdvr_b200/libdvr_b200.so
  echo "== $v ($lib)"
  DVR_LIB_PATH=$lib timeout 300 python tools/bench_configs.py --reps ${REPS:-5} --configs ${CONFIGS:-C1,C2,C3,C4,C5} --out gpurun_out/ab_$v.md > gpurun_out/ab_$v.jsonl 2> gpurun_out/ab_$v.err
  echo "rc=$? $(grep -c config gpurun_out/ab_$v.jsonl) configs"
done
