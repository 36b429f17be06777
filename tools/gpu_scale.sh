#!/bin/bash
# One gpurun --gpus N call: the multi-GPU lines of the round (driver's launch form), into gpurun_out/.
# Usage: bash tools/gpu_scale.sh N [tag]
set -u
N=${1:-2}
TAG=${2:-r01}
OUT=gpurun_out
mkdir -p $OUT
PORT=29511
tr() { # name, script args...
  local name=$1; shift
  PORT=$((PORT+1))
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT "$@" \
      > $OUT/$name.json 2> $OUT/$name.err
  echo "$name rc=$?"; grep '^{' $OUT/$name.json | cut -c1-260
}
tr scale_bench_n${N}_$TAG        bench.py --gpus $N --cpu-seconds 0
tr scale_65536_n${N}_$TAG        bench.py --gpus $N --batch $((65536 / N)) --steps 100 --warmup 10 --cpu-seconds 0
tr scale_update_n${N}_$TAG       bench_update.py --gpus $N --cpu-seconds 0
tr scale_update_wide_n${N}_$TAG  bench_update.py --gpus $N --layers 4 --channels 2048 --steps 20 --warmup 3 --cpu-seconds 0
tr scale_wide_n${N}_$TAG         bench_wide.py --gpus $N --cpu-seconds 0
