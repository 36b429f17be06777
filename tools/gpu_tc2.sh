#!/bin/bash
set -u
OUT=gpurun_out
mkdir -p $OUT
echo "== pytest" ; timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_r02.log 2>&1 ; echo "rc=$?" ; tail -5 $OUT/pytest_r02.log
echo "== trace" ; timeout 300 python tools/trace_phases.py 2>&1 | tee $OUT/r02_trace_phases.txt
for B in 4096 65536; do echo "== quick B=$B"; B=$B timeout 300 python tools/quick_time.py 2>&1 | grep "decide\|encode\|dense" ; done
echo "== bench" ; timeout 900 python bench.py 2>$OUT/bench_r02.err > $OUT/bench_r02.json; python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_r02.json').read())
print({k:d[k] for k in ('value','ms_per_step','gpu_launches')}, d['roofline']['frac'], d['roofline']['launch_us'])
print('e2e', d['e2e']['value'], d['e2e'].get('sync_value'), d['e2e'].get('pageable_input_value'))
print('cpu', d['cpu_baseline'])
for r in d['roofline']['secondary'] or []:
    print({k:(round(v,4) if isinstance(v,float) else v) for k,v in r.items() if k not in ('config',)}, '|', r.get('config','')[:70])
PY
tail -5 $OUT/bench_r02.err
echo "== bench reference"; timeout 600 python bench.py --impl reference --steps 5 --warmup 1 | cut -c1-400; cp gpurun_out/bench_r02.json gpurun_out/bench_n1_r02.json
