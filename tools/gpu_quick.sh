#!/bin/bash
# scratch: one short gpurun call (edit freely); the round's evidence script is tools/gpu_r02.sh
set -u
timeout 400 python -m pytest tests/test_gpu_bn.py -m gpu -q -k "oracle" 2>&1 | tail -30
