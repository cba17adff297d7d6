#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/bench_f64.json 2> gpurun_out/bench_f64.err; echo "bench1 rc=$?"
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 > gpurun_out/bench_f64_n2.json 2> gpurun_out/bench_f64_n2.err; echo "bench2 rc=$?"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --impl reference --steps 1 --warmup 0 > gpurun_out/bench_ref_n2.json 2> gpurun_out/bench_ref_n2.err; echo "ref2 rc=$?"
tail -2 gpurun_out/bench_f64_n2.err
python - <<'PY'
import json
for f in ("bench_f64","bench_f64_n2","bench_ref_n2"):
    try:
        d=json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
        print(f, d["n_gpus"], "%.4g"%d["value"], d.get("roofline",{}).get("step_frac"), d.get("e2e"), d.get("clocks"))
    except Exception as e: print(f, "ERR", e)
PY
