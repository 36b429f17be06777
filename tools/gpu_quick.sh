#!/bin/bash
for S in 4000 0; do echo "== PK_SESSION_SPIN=$S"; PK_SESSION_SPIN=$S timeout 300 python tools/prof_act_batch.py 2>&1 | grep "us per call"; done
echo "== tests"; timeout 900 python -m pytest tests/test_gpu_agents.py tests/test_lockstep.py tests/test_gpu_decide.py -m gpu -x -q 2>&1 | tail -2
timeout 600 python bench.py --steps 20 --warmup 5 --cpu-seconds 0 --no-secondary | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); e=d['e2e']; print('value', d['value'], 'e2e', e['value'], 'pageable', e['pageable_input_value'], 'sync', e['sync_value'], e['sync_ms_per_launch'])"
