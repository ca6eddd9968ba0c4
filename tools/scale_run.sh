#!/bin/bash
# One point of the scaling curves: bench.py on N GPUs of one box for C4 (the headline config) and C5, plus the
# reference arm launched the same way. Usage (through gpurun --gpus N): bash tools/scale_run.sh N [tag]
set -x
N=${1:-2}
TAG=${2:-r02}
mkdir -p gpurun_out
launch() {  # port, then bench.py arguments
  local port=$1; shift
  if [ "$N" = "1" ]; then python bench.py --gpus 1 "$@"
  else python -m torch.distributed.run --nnodes=1 --nproc-per-node "$N" --master-addr 127.0.0.1 --master-port "$port" bench.py --gpus "$N" "$@"
  fi
}
export -f launch; export N
timeout 900 bash -c 'launch 29511 --steps 3 --warmup 3 --no-cpu-baseline' > gpurun_out/${TAG}_scale_C4_n$N.json 2> gpurun_out/${TAG}_scale_C4_n$N.err
timeout 1500 bash -c 'launch 29512 --config C5 --steps 3 --warmup 3 --no-cpu-baseline' > gpurun_out/${TAG}_scale_C5_n$N.json 2> gpurun_out/${TAG}_scale_C5_n$N.err
timeout 600 bash -c 'launch 29513 --impl reference --steps 2 --warmup 1' > gpurun_out/${TAG}_scale_reference_n$N.json 2> gpurun_out/${TAG}_scale_reference_n$N.err
head -c 500 gpurun_out/${TAG}_scale_C4_n$N.json; echo; head -c 500 gpurun_out/${TAG}_scale_C5_n$N.json; echo
head -c 300 gpurun_out/${TAG}_scale_reference_n$N.json; echo; tail -3 gpurun_out/${TAG}_scale_C5_n$N.err
