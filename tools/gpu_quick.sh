#!/bin/bash
# scratch: one short gpurun call (edit freely); the round's evidence script is tools/gpu_r02.sh
set -u
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,memory.total --format=csv > $OUT/gpu_r02.txt 2>&1
echo "== pytest"; timeout 900 python -m pytest tests -m gpu -x -q -s > $OUT/pytest_r02.log 2>&1; echo "pytest rc=$?"; tail -2 $OUT/pytest_r02.log
echo "== smoke"; timeout 300 python __graft_entry__.py --smoke > $OUT/smoke_r02.log 2>&1; echo "smoke rc=$?"; tail -1 $OUT/smoke_r02.log
echo "== bench"; timeout 900 python bench.py --steps 20 --warmup 5 > $OUT/bench_r02.json 2> $OUT/bench_r02.err; echo "bench rc=$?"; cut -c1-200 $OUT/bench_r02.json
echo "== bench reference"; timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > $OUT/bench_ref_r02.json 2> $OUT/bench_ref_r02.err; echo "rc=$?"; cut -c1-200 $OUT/bench_ref_r02.json
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_r02.json'))
print('e2e', d['e2e']['value'], 'frac', d['roofline']['frac'])
for r in d['roofline'].get('secondary',[]):
    print(r.get('leg'), {k:(round(v,3) if isinstance(v,float) else v) for k,v in r.items() if k in ('transitions_per_s','update_ms','decisions_per_s','frac','error','parity_ok')})
PY
