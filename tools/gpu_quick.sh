#!/bin/bash
set -u
OUT=gpurun_out
mkdir -p $OUT
echo "== pytest multi-GPU cases"; timeout 600 python -m pytest tests -m gpu -x -q -k "two_gpus" > $OUT/pytest_multigpu_r02.log 2>&1; echo "rc=$?"; tail -3 $OUT/pytest_multigpu_r02.log
N=2
echo "== bench N=$N"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$N bench.py --gpus $N --steps 20 --warmup 5 > $OUT/scale_bench_n${N}_r02.json 2> $OUT/scale_bench_n${N}_r02.err
echo "rc=$?"; tail -2 $OUT/scale_bench_n${N}_r02.err
N=$N python - <<'PY'
import json,os
n=os.environ['N']
d=json.loads(open(f'gpurun_out/scale_bench_n{n}_r02.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','ms_per_step','gpu_launches','n_gpus')}, d['roofline']['frac'], 'e2e', d['e2e']['value'])
for r in d['roofline']['secondary'] or []:
    print({k:(round(v,4) if isinstance(v,float) else v) for k,v in r.items() if k in ('leg','exchange','decisions_per_s','transitions_per_s','frac','parity_ok','error')})
PY
