#!/bin/bash
set -u
for M in 1 2; do
echo "== MINB=$M"; PK_DECIDE_MINB=$M timeout 600 python bench.py --no-secondary --cpu-seconds 0 --no-e2e --steps 60 | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print(d['value'], d['roofline']['frac'], d['roofline']['launch_us'])"
B=65536 PK_PDL=1 PK_DECIDE_MINB=$M timeout 300 python tools/quick_time.py 2>&1 | grep "decide+feat"
done
PK_DECIDE_MINB=2 timeout 300 python -m pytest tests/test_gpu_decide.py -m gpu -x -q 2>&1 | tail -2
