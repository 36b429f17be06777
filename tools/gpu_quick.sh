#!/bin/bash
set -u
OUT=gpurun_out
mkdir -p $OUT
echo "== pytest decide"; timeout 600 python -m pytest tests/test_gpu_decide.py tests/test_gpu_agents.py -m gpu -x -q 2>&1 | tail -3
echo "== trace"; timeout 300 python tools/trace_phases.py 2>&1 | head -11
for B in 4096 65536; do for S in 0 1; do echo "== quick B=$B STAGE=$S"; B=$B PK_PDL=1 PK_STAGE_STATE=$S timeout 300 python tools/quick_time.py 2>&1 | grep "decide+feat\|encode"; done; done
echo "== bench"; timeout 600 python bench.py --no-secondary --cpu-seconds 0 --steps 100 | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print(d['value'], d['roofline']['frac'], d['roofline']['launch_us'], 'e2e', d['e2e']['value'])"
