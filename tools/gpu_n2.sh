#!/bin/bash
# two-GPU checks: the 2-GPU pytest cases and bench.py at N = 2 (secondary legs: both exchange routes + parity_ok)
set -u
OUT=gpurun_out
mkdir -p $OUT
echo "== pytest (2 GPUs visible)"; timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_n2_r02.log 2>&1; echo "rc=$?"; tail -4 $OUT/pytest_n2_r02.log
echo "== bench N=2"
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 > $OUT/bench_n2_r02.json 2> $OUT/bench_n2_r02.err
echo "rc=$?"; tail -3 $OUT/bench_n2_r02.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_n2_r02.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','ms_per_step','gpu_launches','n_gpus')}, d['roofline']['frac'], d['roofline']['launch_us'])
print('e2e', d['e2e']['value'])
for r in d['roofline']['secondary'] or []:
    print({k:(round(v,4) if isinstance(v,float) else v) for k,v in r.items() if k not in ('config',)}, '|', r.get('config','')[:60])
PY
