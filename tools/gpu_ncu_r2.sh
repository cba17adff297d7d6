#!/bin/bash
# ncu evidence of round 2 (one GPU).  Every profiled command first runs to completion WITHOUT ncu.
set -u
mkdir -p gpurun_out
O=gpurun_out
K='regex:pipe_kernel|scan_|ensemble_|setup_|kf_forward_kernel|reduce_rows|pack_rows|unpack_rows'
# (1) launch list of the bench command (short: 2 timed steps), all of this library's kernels
CMD="python bench.py --steps 2 --warmup 3 --no-cpu --cpu-seconds 1"
$CMD > $O/ncu_plain1.log 2>&1 && \
timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -k "$K" --csv --log-file $O/launches.csv $CMD > $O/ncu1.log 2>&1
echo "ncu launches rc=$?"
# (2) DRAM bytes at full size (1 pass, no replay) of the two dominant kernels of the timed step
CMD2="python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e --no-c5 --no-setup --no-sustained --no-grid --no-parity"
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:pipe_kernel -s 6 -c 4 --csv --log-file $O/dram_full.csv $CMD2 > $O/ncu2.log 2>&1
echo "ncu dram rc=$?"
# (2b) the same for f32 (one 16384-track launch per pass)
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:pipe_kernel -s 6 -c 4 --csv --log-file $O/dram_full_f32.csv $CMD2 --dtype f32 > $O/ncu2b.log 2>&1
echo "ncu dram f32 rc=$?"
# (2c) DRAM bytes of every kernel of one pca_setup call at the bench's group shape (8192 x 100000, f64)
python tools/setup_time.py bench > $O/ncu_plain2c.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:'setup_|center_rows' --csv --log-file $O/dram_setup.csv python tools/setup_time.py bench > $O/ncu2c.log 2>&1
echo "ncu setup rc=$?"
# (3) DRAM bytes of configs[4] (64 x 10 M): every kernel of one routed forward + backward call
CMD3="python tools/scan_time.py"
SCAN_TIME_BIG_ONLY=1 SCAN_TIME_F64_ONLY=1 $CMD3 > $O/ncu_plain3.log 2>&1 && \
SCAN_TIME_BIG_ONLY=1 SCAN_TIME_F64_ONLY=1 timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:'pipe_kernel|scan_' -c 40 --csv --log-file $O/dram_c5.csv $CMD3 > $O/ncu3.log 2>&1
echo "ncu c5 rc=$?"
# (4) --set full of the dominant kernels at T = 20000 (source-level view)
CMD4="python bench.py --steps 2 --warmup 3 --no-cpu --no-e2e --no-c5 --no-setup --no-sustained --no-grid --no-parity --frames 20000"
$CMD4 > $O/ncu_plain4.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:pipe_kernel -s 6 -c 2 -f -o $O/prof_pipe_r2 $CMD4 > $O/ncu4.log 2>&1
echo "ncu full rc=$?"
