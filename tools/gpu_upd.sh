#!/bin/bash
set -u
OUT=gpurun_out
mkdir -p $OUT
echo "== pytest update"; timeout 900 python -m pytest tests/test_gpu_update.py tests/test_lockstep.py tests/test_gpu_agents.py -m gpu -x -q 2>&1 | tail -4
for F in 1 0; do echo "== bench_update FUSED=$F"; PK_UPDATE_FUSED=$F timeout 600 python bench_update.py --cpu-seconds 0 --steps 50 | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'])"; done
echo "== launches"; timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file $OUT/r02_launches_update.csv python bench_update.py --cpu-seconds 0 --steps 3 --warmup 3 > /dev/null 2>&1; python - <<'PY'
import csv
rows=[r for r in csv.reader(open('gpurun_out/r02_launches_update.csv')) if len(r)>10 and r[0].isdigit()]
for r in rows[-12:]: print(r[4][:60], r[-1])
PY
