#!/bin/bash
set -u
mkdir -p gpurun_out
echo "== pytest dense tc"; timeout 300 python -m pytest tests/test_gpu_decide.py -m gpu -x -q -k "dense" 2>&1 | tail -15
for B in 4096 65536; do for TC in 1 0; do
echo "== quick B=$B PK_DENSE_TC=$TC"; B=$B PK_DENSE_TC=$TC timeout 120 python tools/quick_time.py 2>&1 | grep "^dense\|^B="
done; done
PK_DENSE_TC=1 B=65536 timeout 120 python tools/trace_dense.py 2>&1 | tail -60
