#!/bin/bash
# ncu --set full of the scan summary kernels (f64, 64 tracks x 10 M frames), one launch each; run only after the
# same command has exited 0 without ncu
mkdir -p gpurun_out
SCAN_TIME_BIG_ONLY=1 SCAN_TIME_F64_ONLY=1 python tools/scan_time.py > gpurun_out/scan_plain.log 2>&1 || exit 1
SCAN_TIME_BIG_ONLY=1 SCAN_TIME_F64_ONLY=1 timeout 1200 ncu --set full --clock-control none --import-source on \
  -k regex:scan_.*summary -s 2 -c 2 -f -o gpurun_out/prof_scan python tools/scan_time.py > gpurun_out/scan_ncu_full.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/scan_plain.log
