#!/bin/bash
# One GPU-box session: tests, smoke, bench (both dtypes + reference arm), ncu evidence.
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/rc.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/rc.txt
timeout 600 python bench.py > gpurun_out/bench_f64.json 2> gpurun_out/bench_f64.err; echo "bench f64 rc=$?" | tee -a gpurun_out/rc.txt
timeout 600 python bench.py --dtype f32 --no-cpu > gpurun_out/bench_f32.json 2> gpurun_out/bench_f32.err; echo "bench f32 rc=$?" | tee -a gpurun_out/rc.txt
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "bench ref rc=$?" | tee -a gpurun_out/rc.txt
timeout 600 python tools/scan_time.py > gpurun_out/scan_time.log 2>&1; echo "scan rc=$?" | tee -a gpurun_out/rc.txt
timeout 300 python tools/setup_time.py > gpurun_out/setup_time.log 2>&1; echo "setup rc=$?" | tee -a gpurun_out/rc.txt
# ncu: launch list of the bench command, DRAM bytes at full size (1 pass, no replay), --set full at T=20000
CMD="python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e"
$CMD > gpurun_out/plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:kernel --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu1.log 2>&1
echo "ncu launches rc=$?" | tee -a gpurun_out/rc.txt
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:pipe_kernel -s 6 -c 4 --csv --log-file gpurun_out/dram_full.csv $CMD > gpurun_out/ncu2.log 2>&1
echo "ncu dram rc=$?" | tee -a gpurun_out/rc.txt
CMD2="python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e --frames 20000"
$CMD2 > gpurun_out/plain2.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:pipe_kernel -s 6 -c 2 -f -o gpurun_out/prof_pipe $CMD2 > gpurun_out/ncu3.log 2>&1
echo "ncu full rc=$?" | tee -a gpurun_out/rc.txt
cat gpurun_out/rc.txt
