#!/bin/bash
set -u
OUT=gpurun_out
mkdir -p $OUT
N=8
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29538 bench.py --gpus $N --steps 5 --warmup 3 --no-e2e --skip-legs shard65536,wide,wide_update,gpu_reference,concurrent64 > $OUT/b8.json 2> $OUT/b8.err
echo "rc=$?"; tail -3 $OUT/b8.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/b8.json').read().strip().splitlines()[-1])
for r in d['roofline']['secondary'] or []:
    print({k:(round(v,4) if isinstance(v,float) else v) for k,v in r.items() if k in ('leg','exchange','transitions_per_s','ms_per_step','parity_ok','parity','error')})
PY
echo "== launch list of one rank's update step (nsys-free: CUDA events around each kernel are not available; ncu on rank 0 only is not possible under torchrun)"
