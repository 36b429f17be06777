#!/bin/bash
# kernel iteration on pk_decide_tc_kernel: parity tests, timings with and without it, phase trace, bench
set -u
OUT=gpurun_out
mkdir -p $OUT
echo "== pytest decide" ; timeout 600 python -m pytest tests/test_gpu_decide.py -m gpu -x -q > $OUT/pytest_tc.log 2>&1 ; echo "rc=$?" ; tail -15 $OUT/pytest_tc.log
for B in 4096 65536; do
  for TC in 0 1; do echo "== quick B=$B TC=$TC"; B=$B PK_DECIDE_TC=$TC timeout 300 python tools/quick_time.py 2>&1 | grep -v "^dense\|^memset" ; done
done
echo "== trace" ; timeout 300 python tools/trace_tc.py 2>&1 | tee $OUT/trace_tc.txt
echo "== bench" ; timeout 600 python bench.py --cpu-seconds 0 2>&1 | tee $OUT/bench_tc.json | cut -c1-1200
