#!/bin/bash
set -u
timeout 400 python -m pytest tests/test_gpu_bn.py -m gpu -q 2>&1 | tail -4
timeout 200 python tools/prof_bn_update.py 2>&1 | tail -2
