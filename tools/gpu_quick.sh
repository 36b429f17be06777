#!/bin/bash
set -u
echo "== pytest dense, PK_DENSE_2CTA=1"; PK_DENSE_2CTA=1 timeout 120 python -m pytest tests/test_gpu_decide.py -m gpu -x -q -k "dense" 2>&1 | tail -8
echo "rc=$?"
for C in 1 0; do for B in 4096 65536; do echo "== PK_DENSE_2CTA=$C B=$B"; PK_DENSE_2CTA=$C B=$B timeout 90 python tools/quick_time.py 2>&1 | grep "^dense"; done; done
