#!/bin/bash
set -u
timeout 400 python -m pytest tests/test_gpu_bn.py -m gpu -q 2>&1 | tail -30
timeout 300 python bench.py --steps 5 --warmup 3 --cpu-seconds 0 --skip-legs shard65536,dense,update,wide,wide_update,gpu_reference,concurrent64 > gpurun_out/bench_bn.json 2> gpurun_out/bench_bn.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_bn.json'))
for r in d['roofline'].get('secondary',[]):
    print(r.get('leg'), {k:(round(v,3) if isinstance(v,float) else v) for k,v in r.items() if k in ('transitions_per_s','update_ms','error','loss_finite','leg_wall_s')})
PY
