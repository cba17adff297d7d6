#!/bin/bash
# Developer probe: end-to-end (host buffers) throughput of batched.smooth_tracks vs its block size.
set -u
mkdir -p gpurun_out
for BLK in ${BLOCKS:-512 384 256 128}; do
  timeout 600 python bench.py --tracks 2048 --steps 3 --warmup 3 --no-c5 --no-grid --no-parity --no-cpu --no-setup --no-sustained --e2e-block $BLK \
    2> gpurun_out/e2e_blk$BLK.err | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1]); e = d['e2e']
print('block $BLK', round(e['value'] / 1e9, 4), 'G', round(e['ms_per_step'], 1), 'ms', 'pcie_frac', round(e['pcie_frac'], 3), e['pcie'])
" | tee -a gpurun_out/e2e_blocks.txt
done
