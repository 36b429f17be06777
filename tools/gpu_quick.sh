#!/bin/bash
# scratch: one short gpurun call (edit freely); the round's evidence script is tools/gpu_r02.sh
set -u
OUT=gpurun_out
mkdir -p $OUT
echo "== ncu launches"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches_r02.csv \
    python bench.py --steps 2 --warmup 3 --cpu-seconds 0 --no-graph --no-secondary --no-e2e > $OUT/ncu_launches_r02.log 2>&1
echo "ncu launches rc=$?"
echo "== ncu full"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:pk_decide_kernel -s 4 -c 2 -f -o $OUT/prof_r02 \
    python tools/prof_decide.py > $OUT/ncu_full_r02.log 2>&1
echo "ncu full rc=$?"
ncu -i $OUT/prof_r02.ncu-rep --page raw --csv > $OUT/r02_ncu_full_raw.csv 2>/dev/null
ls -la $OUT | grep -i "launches_r02\|prof_r02\|raw"
