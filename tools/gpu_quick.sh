#!/bin/bash
set -u
OUT=gpurun_out
mkdir -p $OUT
echo "== pytest update"; timeout 900 python -m pytest tests/test_gpu_update.py tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -3
for NB in 32768 4096; do echo "== bench_update --batch $NB"; timeout 300 python bench_update.py --cpu-seconds 0 --steps 100 --batch $NB | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'])"; done
echo "== launches"; timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file $OUT/r02_launches_update.csv python bench_update.py --cpu-seconds 0 --steps 3 --warmup 3 > /dev/null 2>&1; python - <<'PY'
import csv
rows=[r for r in csv.reader(open('gpurun_out/r02_launches_update.csv')) if len(r)>10 and r[0].isdigit()]
for r in rows[-8:]: print(r[4][:60], r[-1])
PY
