#!/bin/bash
# Developer probe: time every tuning build of the library (eks_b200/libeksb*.so) on the same shapes.
# usage: tools/tune.sh [bits:B:T ...]   (REPS env = repetitions; sustained runs show power-cap behaviour)
mkdir -p gpurun_out
: > gpurun_out/tune.log
SHAPES="${@:-64:8192:20000 32:16384:20000 64:16384:10000 32:8192:20000}"
for lib in eks_b200/libeksb*.so; do
  echo "== $lib" >> gpurun_out/tune.log
  EKSB_LIBRARY=$PWD/$lib timeout 300 python tools/quick_time.py $SHAPES >> gpurun_out/tune.log 2>&1
done
cat gpurun_out/tune.log
