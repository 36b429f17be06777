#!/bin/bash
# One gpurun call: every single-GPU bench line of the round, written to gpurun_out/ under the names profiles/ uses.
# Usage (repo root on the GPU box): bash tools/gpu_benches.sh [tag]
set -u
TAG=${1:-r01}
OUT=gpurun_out
mkdir -p $OUT
run() { # name, command...
  local name=$1; shift
  timeout 600 "$@" > $OUT/$name.json 2> $OUT/$name.err
  echo "$name rc=$?"; cut -c1-400 $OUT/$name.json
}
run bench_$TAG                 python bench.py
run bench_65536_n1_$TAG        python bench.py --batch 65536 --steps 100 --warmup 10 --cpu-seconds 0
run bench_b64_$TAG             python bench.py --batch 64 --cpu-seconds 0
run bench_update_n1_$TAG       python bench_update.py
run bench_update_wide_n1_$TAG  python bench_update.py --layers 4 --channels 2048 --steps 20 --warmup 3 --cpu-seconds 0
run bench_wide_n1_$TAG         python bench_wide.py
run bench_wide_train_n1_$TAG   python bench_wide.py --mode train
timeout 300 python tools/bench_party.py > $OUT/party_$TAG.log 2>&1; echo "party rc=$?"; tail -5 $OUT/party_$TAG.log
