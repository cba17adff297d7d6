#!/bin/bash
# grouped pipelined launches: tests, then a sweep of the group size at large track counts
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests.log 2>&1; tail -4 gpurun_out/gpu_tests.log
for g in 148 256 296 444 512 592; do echo "f64 group $g"; EKSB_PIPE_MAX_TILES=$g EKSB_PIPE_GROUP_TILES=$g python tools/quick_time.py 64:65536:8000 64:20000:8000 2>&1 | sed -n '2p;4p'; done
for g in 296 512 592 888 1184; do echo "f32 group $g"; EKSB_PIPE_MAX_TILES=$g EKSB_PIPE_GROUP_TILES=$g python tools/quick_time.py 32:131072:8000 32:40000:8000 2>&1 | sed -n '2p;4p'; done
echo "single launch (old policy below 1536 tiles)"; EKSB_PIPE_MAX_TILES=1536 python tools/quick_time.py 64:20000:8000 64:32768:8000 32:40000:8000 2>&1 | sed -n '2p;4p;6p'
