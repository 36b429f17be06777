#!/bin/bash
set -u
timeout 400 python -m pytest tests/test_gpu_bn.py -m gpu -q 2>&1 | tail -40
