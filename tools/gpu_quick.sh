#!/bin/bash
set -u
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29583 tools/check_wide_overlap.py 2>&1 | grep -v "OMP_NUM\|\*\*\*\*" | grep -B2 -A14 "Traceback" | head -60
