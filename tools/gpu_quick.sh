#!/bin/bash
set -u
echo "== pytest dense tc"; timeout 200 python -m pytest tests/test_gpu_decide.py tests/test_gpu_agents.py -m gpu -x -q -k "dense" 2>&1 | tail -5
for B in 4096 65536; do B=$B timeout 120 python tools/quick_time.py 2>&1 | grep "^dense"; done
