#!/bin/bash
set -u
timeout 200 python tools/time_update_shard.py 2>&1 | tail -6
SHARDS=512 ITERS=3 timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/upd512.csv python tools/time_update_shard.py > /dev/null 2>&1
SHARDS=4096 ITERS=3 timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/upd4096.csv python tools/time_update_shard.py > /dev/null 2>&1
tail -8 gpurun_out/upd512.csv | cut -d, -f5,12-15 | cut -c1-200
tail -8 gpurun_out/upd4096.csv | cut -d, -f5,12-15 | cut -c1-200
