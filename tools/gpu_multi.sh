#!/bin/bash
# Multi-GPU session (gpurun --gpus N -- bash tools/gpu_multi.sh N): the NCCL grid-search test and the bench at N ranks.
set -u
N=${1:-2}
mkdir -p gpurun_out
O=gpurun_out
nvidia-smi --query-gpu=index,name,memory.total --format=csv > $O/smi_n$N.txt 2>&1
nvidia-smi topo -m >> $O/smi_n$N.txt 2>&1
if [ "$N" = "2" ]; then
  timeout 600 python -m pytest tests/test_gpu_properties.py -m gpu -x -q -k "nccl" > $O/pytest_nccl.log 2>&1; echo "pytest nccl rc=$?"; tail -3 $O/pytest_nccl.log
  timeout 600 python -m pytest tests/test_gpu_round2.py -m gpu -x -q -k "device_lanes" > $O/pytest_lanes.log 2>&1; echo "pytest lanes rc=$?"; tail -3 $O/pytest_lanes.log
fi
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 > $O/bench_f64_n$N.json 2> $O/bench_f64_n$N.err; echo "bench N=$N rc=$?"
tail -c 600 $O/bench_f64_n$N.err
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus $N --steps 1 --warmup 1 > $O/bench_ref_n$N.json 2> $O/bench_ref_n$N.err; echo "bench ref N=$N rc=$?"
