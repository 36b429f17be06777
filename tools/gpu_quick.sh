#!/bin/bash
for W in 13 14 15 16 18 20; do echo "== PK_DECIDE_WARPS=$W"; PK_DECIDE_WARPS=$W timeout 300 python tools/quick_streams.py 2>&1 | grep "pdl=1 streams=1"; done
