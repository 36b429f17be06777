#!/bin/bash
set -u
python -m pokeai_b200.build | tail -1
echo "== pytest lockstep/protocol"; timeout 600 python -m pytest tests/test_lockstep.py tests/test_protocol.py -x -q 2>&1 | tail -3
for P in 0 1; do echo "== PK_PY_PACKER=$P"; PK_PY_PACKER=$P timeout 600 python bench.py --steps 5 --warmup 3 --cpu-seconds 0 --no-e2e --skip-legs shard65536,dense,update,wide_update,gpu_reference,wide | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
for r in d['roofline']['secondary']: print({k:(round(v,2) if isinstance(v,float) else v) for k,v in r.items() if k in ('leg','decisions_per_s','round_trip_us','host_pack_us_per_round','gpu_call_us_per_round','cpu_reference_round_us','error')})"
done
