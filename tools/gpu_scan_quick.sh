#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -k "scan or parallel_in_time or block_tridiagonal" > gpurun_out/scan_tests.log 2>&1; tail -3 gpurun_out/scan_tests.log
SCAN_TIME_BIG_ONLY=1 python tools/scan_time.py > gpurun_out/scan_new.log 2>&1; cat gpurun_out/scan_new.log
for w in 8 16 24; do echo "warps/SM $w"; EKSB_SCAN_WARPS_PER_SM=$w SCAN_TIME_BIG_ONLY=1 python tools/scan_time.py 2>&1 | grep "chunk" | sed -n "2p;4p"; done
