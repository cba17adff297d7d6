#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -k "setup or select or session or multi_cam or groups" > gpurun_out/setup_tests.log 2>&1; tail -5 gpurun_out/setup_tests.log
python tools/setup_time.py > gpurun_out/setup_time.log 2>&1; cat gpurun_out/setup_time.log
echo "full-length passes"; EKSB_SETUP_NO_COLLECT=1 python tools/setup_time.py 2>&1 | tail -9
