#!/bin/bash
# One gpurun call: GPU tests, smoke, bench (both arms), kernel timing sweep, ncu launch list + full capture.
# Usage (from the repo root on the GPU box): bash tools/gpu_round.sh [tag]
set -u
TAG=${1:-r01}
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,memory.total --format=csv > $OUT/gpu_$TAG.txt 2>&1
echo "== pytest" ; timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_$TAG.log 2>&1 ; echo "pytest rc=$?" ; tail -5 $OUT/pytest_$TAG.log
echo "== smoke" ; timeout 300 python __graft_entry__.py --smoke > $OUT/smoke_$TAG.log 2>&1 ; echo "smoke rc=$?" ; tail -2 $OUT/smoke_$TAG.log
echo "== bench" ; timeout 600 python bench.py > $OUT/bench_$TAG.json 2> $OUT/bench_$TAG.err ; echo "bench rc=$?" ; cat $OUT/bench_$TAG.json ; tail -3 $OUT/bench_$TAG.err
echo "== bench reference" ; timeout 600 python bench.py --impl reference --steps 10 --warmup 2 > $OUT/bench_ref_$TAG.json 2> $OUT/bench_ref_$TAG.err ; echo "rc=$?" ; cut -c1-300 $OUT/bench_ref_$TAG.json
echo "== quick_time" ; for B in 64 4096 65536; do B=$B timeout 300 python tools/quick_time.py ; done > $OUT/quick_$TAG.log 2>&1 ; cat $OUT/quick_$TAG.log
echo "== ncu launches"
timeout 300 python bench.py --steps 5 --warmup 3 --cpu-seconds 0 > $OUT/plain_$TAG.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches_$TAG.csv \
    python bench.py --steps 5 --warmup 3 --cpu-seconds 0 > $OUT/ncu_launches_$TAG.log 2>&1
echo "ncu launches rc=$?"
echo "== ncu full"
timeout 300 python tools/prof_decide.py > $OUT/plain_prof_$TAG.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:pk_decide -s 6 -c 3 -f -o $OUT/prof_$TAG \
    python tools/prof_decide.py > $OUT/ncu_full_$TAG.log 2>&1
echo "ncu full rc=$?"
ls -la $OUT
