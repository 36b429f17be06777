#!/bin/bash
set -u
echo "== pytest wide"; timeout 600 python -m pytest tests/test_gpu_wide.py tests/test_gpu_agents.py -m gpu -x -q 2>&1 | tail -3
for F in 1 0; do echo "== PK_WIDE_FUSED_HEAD=$F"; PK_WIDE_FUSED_HEAD=$F timeout 600 python bench.py --steps 5 --warmup 3 --cpu-seconds 0 --no-e2e --skip-legs shard65536,dense,update,wide_update,gpu_reference,concurrent64 | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
for r in d['roofline']['secondary']: print({k:(round(v,4) if isinstance(v,float) else v) for k,v in r.items() if k in ('leg','per_s','decisions_per_s','ms_per_step','frac_burst','frac_sustained','error','tflops')})"
done
