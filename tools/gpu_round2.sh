#!/bin/bash
# One GPU-box session of round 2: tests, smoke, bench (both dtypes + reference arm), shard-sized runs.
set -u
mkdir -p gpurun_out
O=gpurun_out
rm -f $O/rc2.txt
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw,memory.total --format=csv > $O/smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/rc2.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $O/rc2.txt
timeout 900 python bench.py > $O/bench_f64.json 2> $O/bench_f64.err; echo "bench f64 rc=$?" | tee -a $O/rc2.txt
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err; echo "bench ref rc=$?" | tee -a $O/rc2.txt
timeout 900 python bench.py --dtype f32 --no-cpu > $O/bench_f32.json 2> $O/bench_f32.err; echo "bench f32 rc=$?" | tee -a $O/rc2.txt
# what ONE GPU of an N-GPU run does with its shard of the fixed 16384-track problem (N = 4: 4096, N = 8: 2048 tracks)
for S in 4096 2048; do
  timeout 600 python bench.py --tracks $S --no-cpu --no-e2e --no-c5 --no-setup --no-sustained > $O/bench_f64_shard$S.json 2> $O/bench_f64_shard$S.err; echo "bench shard $S rc=$?" | tee -a $O/rc2.txt
done
ITERS=200 timeout 600 python tools/stress_parity.py > $O/stress_parity.log 2>&1; echo "stress rc=$?" | tee -a $O/rc2.txt
cat $O/rc2.txt
