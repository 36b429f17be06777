#!/bin/bash
set -u
mkdir -p gpurun_out
echo "== pytest wide"; timeout 300 python -m pytest tests/test_gpu_wide.py -m gpu -x -q 2>&1 | tail -3
timeout 600 python bench.py --steps 10 --warmup 3 --cpu-seconds 0 --no-e2e --skip-legs shard65536,dense,update,gpu_reference,concurrent64 | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
for r in d['roofline']['secondary']: print({k:(round(v,4) if isinstance(v,float) else v) for k,v in r.items() if k in ('leg','per_s','transitions_per_s','ms_per_step','frac_burst','error')})"
echo "== ncu wide launches"; timeout 600 ncu --metrics gpu__time_duration.sum,sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active --clock-control none -k regex:pk_wide_gemm -c 40 --csv --log-file gpurun_out/r02_launches_wide.csv python bench_wide.py --steps 2 --warmup 3 --cpu-seconds 0 --mode train > /dev/null 2>&1
python - <<'PY'
import csv,collections
rows=[r for r in csv.reader(open('gpurun_out/r02_launches_wide.csv')) if len(r)>10 and r[0].isdigit()]
d=collections.OrderedDict()
for r in rows: d.setdefault((r[0]),{'k':r[4][:44]})[r[-3][:30]]=r[-1]
for k,v in list(d.items())[-11:]: print(k, v)
PY
