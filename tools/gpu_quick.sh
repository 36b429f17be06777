#!/bin/bash
set -u
mkdir -p gpurun_out
echo "== 8-GPU check"; timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29583 tools/check_wide_overlap.py 2>&1 | grep "wide "
echo "== bench N=8 wide_update"; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29528 bench.py --gpus 8 --steps 5 --warmup 3 --no-e2e --skip-legs shard65536,dense,update,wide,gpu_reference,concurrent64 > gpurun_out/b.json 2> gpurun_out/b.err; echo "rc=$?"; tail -2 gpurun_out/b.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/b.json').read().strip().splitlines()[-1])
for r in d['roofline']['secondary'] or []:
    print({k:(round(v,4) if isinstance(v,float) else v) for k,v in r.items() if k in ('leg','transitions_per_s','ms_per_step','parity_ok','parity','error','frac_burst')}, (r.get('exchange') or '')[:40])
PY
