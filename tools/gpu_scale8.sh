#!/bin/bash
set -u
N=${1:-8}
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus $N --no-cpu > gpurun_out/bench_f64_n$N.json 2> gpurun_out/bench_f64_n$N.err; echo "bench$N rc=$?"
python - <<PY
import json
d=json.loads(open("gpurun_out/bench_f64_n$N.json").read().strip().splitlines()[-1])
print(d["n_gpus"], "%.4g"%d["value"], d["ms_per_step"], d.get("roofline",{}).get("step_frac"), d.get("e2e",{}).get("value"), d.get("clocks"))
PY
