#!/bin/bash
set -u
OUT=gpurun_out
mkdir -p $OUT
echo "== pytest wide+decide"; timeout 600 python -m pytest tests/test_gpu_wide.py tests/test_gpu_decide.py -m gpu -x -q 2>&1 | tail -3
echo "== trace TC"; PK_DECIDE_TC=1 timeout 300 python tools/trace_phases.py 2>&1 | head -11
for B in 4096 65536; do for T in 0 1; do echo "== quick B=$B TC=$T"; B=$B PK_PDL=1 PK_DECIDE_TC=$T timeout 300 python tools/quick_time.py 2>&1 | grep "decide+feat"; done; done
echo "== ncu launches (headline only)"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches_r02.csv \
    python bench.py --steps 2 --warmup 3 --cpu-seconds 0 --no-graph --no-secondary --no-e2e > $OUT/ncu_launches_r02.log 2>&1
echo "rc=$?"
