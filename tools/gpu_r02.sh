#!/bin/bash
# One gpurun call for the round-2 evidence: GPU tests, smoke, bench (both arms), kernel timings, phase trace,
# ncu launch list of bench.py and one full capture of the decision kernel.  Usage: bash tools/gpu_r02.sh
set -u
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,memory.total --format=csv > $OUT/gpu_r02.txt 2>&1
echo "== pytest"; timeout 1200 python -m pytest tests -m gpu -x -q -s > $OUT/pytest_r02.log 2>&1; echo "pytest rc=$?"; tail -3 $OUT/pytest_r02.log
echo "== smoke"; timeout 300 python __graft_entry__.py --smoke > $OUT/smoke_r02.log 2>&1; echo "smoke rc=$?"; tail -1 $OUT/smoke_r02.log
echo "== bench"; timeout 900 python bench.py --steps 20 --warmup 5 > $OUT/bench_r02.json 2> $OUT/bench_r02.err; echo "bench rc=$?"; cut -c1-400 $OUT/bench_r02.json
echo "== bench reference"; timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > $OUT/bench_ref_r02.json 2> $OUT/bench_ref_r02.err; echo "rc=$?"; cut -c1-300 $OUT/bench_ref_r02.json
echo "== quick"; for B in 64 4096 65536; do B=$B PK_PDL=1 timeout 300 python tools/quick_time.py; done > $OUT/quick_r02.log 2>&1; grep -v memset $OUT/quick_r02.log
echo "== quick TC"; B=4096 PK_PDL=1 PK_DECIDE_TC=1 timeout 300 python tools/quick_time.py 2>&1 | grep "decide+feat" | tee $OUT/quick_tc_r02.log
echo "== trace"; timeout 300 python tools/trace_phases.py > $OUT/r02_trace_phases.txt 2>&1; head -11 $OUT/r02_trace_phases.txt
echo "== ncu launches"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches_r02.csv \
    python bench.py --steps 2 --warmup 3 --cpu-seconds 0 --no-graph --no-secondary --no-e2e > $OUT/ncu_launches_r02.log 2>&1
echo "ncu launches rc=$?"
echo "== ncu full"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:pk_decide_kernel -s 4 -c 2 -f -o $OUT/prof_r02 \
    python tools/prof_decide.py > $OUT/ncu_full_r02.log 2>&1
echo "ncu full rc=$?"
ncu -i $OUT/prof_r02.ncu-rep --page raw --csv > $OUT/r02_ncu_full_raw.csv 2>/dev/null
echo "== dense forward: timeline of the roles, ncu full capture"
B=65536 IT0=20 IT1=44 timeout 300 python tools/trace_dense.py > $OUT/r02_trace_dense.txt 2>&1; tail -2 $OUT/r02_trace_dense.txt
timeout 900 ncu --set full --clock-control none --import-source on -k regex:pk_dense_tc_kernel -s 2 -c 1 -f -o $OUT/prof_dense_r02 \
    python tools/prof_dense.py > $OUT/ncu_dense_r02.log 2>&1
echo "ncu dense rc=$?"
ls -la $OUT | head -40
