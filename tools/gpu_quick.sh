#!/bin/bash
set -u
for E in 0 1; do echo "== PK_EARLY_FILL=$E"; PK_EARLY_FILL=$E timeout 300 python -m pytest tests/test_gpu_decide.py -m gpu -x -q 2>&1 | tail -1
for B in 4096 65536; do PK_EARLY_FILL=$E PK_PDL=1 B=$B timeout 120 python tools/quick_time.py 2>&1 | grep "^decide+feat"; done
PK_EARLY_FILL=$E timeout 300 python bench.py --steps 20 --warmup 5 --cpu-seconds 0 --no-secondary --no-e2e | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('bench', d['value'], d['ms_per_step']/16*1e3, 'us/launch frac', d['roofline']['frac'])"
done
