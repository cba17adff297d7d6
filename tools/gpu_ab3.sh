#!/bin/bash
# A/B of the split-chain kernels (forward three-role, backward cov/mean) against the two-role ones + parity tests.
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee gpurun_out/rc.txt
tail -3 gpurun_out/pytest_gpu.log
SH="${SH:-64:4736:100000 32:4736:100000 64:2048:20000 32:2048:20000 64:64:3000}"
for rep in 1 2; do
  echo "== split chains (<= 1 tile per SM)"; REPS=4 timeout 300 python tools/quick_time.py $SH 2>&1 | awk '{print $1,$2,$3,$5,$8,$14,$15,$18,$20}'
  echo "== two-role"; EKSB_BWD_NO_SOLO=1 EKSB_BWD_NO_SPLIT=1 EKSB_FWD3_MAX=0 REPS=4 timeout 300 python tools/quick_time.py $SH 2>&1 | awk '{print $1,$2,$3,$5,$8,$14,$15,$18,$20}'
done 2>&1 | tee gpurun_out/ab3.log
