#!/bin/bash
# full GPU test suite, then the scan timing probe (config 5 shape)
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests.log 2>&1; tail -4 gpurun_out/gpu_tests.log
SCAN_TIME_BIG_ONLY=1 python tools/scan_time.py > gpurun_out/scan_new.log 2>&1; cat gpurun_out/scan_new.log
