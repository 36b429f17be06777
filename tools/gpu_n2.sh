#!/bin/bash
set -u
OUT=gpurun_out
mkdir -p $OUT
for N in 1 2; do
echo "== bench N=$N (update leg only)"
if [ $N -eq 1 ]; then timeout 600 python bench.py --steps 5 --warmup 3 --cpu-seconds 0 --no-e2e --skip-legs shard65536,wide,wide_update,gpu_reference,concurrent64 > $OUT/b.json 2> $OUT/b.err
else timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2952$N bench.py --gpus $N --steps 5 --warmup 3 --no-e2e --skip-legs shard65536,wide,wide_update,gpu_reference,concurrent64 > $OUT/b.json 2> $OUT/b.err; fi
echo "rc=$?"; tail -3 $OUT/b.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/b.json').read().strip().splitlines()[-1])
for r in d['roofline']['secondary'] or []:
    print({k:(round(v,4) if isinstance(v,float) else v) for k,v in r.items() if k in ('leg','exchange','transitions_per_s','ms_per_step','parity_ok','parity','error')})
PY
done
echo "== pytest update"; timeout 900 python -m pytest tests/test_gpu_update.py -m gpu -x -q 2>&1 | tail -3
