#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests.log 2>&1; tail -3 gpurun_out/gpu_tests.log
python tools/quick_time.py 64:8192:20000 64:4736:20000 32:16384:20000 64:65536:8000 2>&1 | sed -n '2p;4p;6p;8p'
timeout 600 python bench.py --no-cpu > gpurun_out/bench_f64.json 2> gpurun_out/bench_f64.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open("gpurun_out/bench_f64.json").read().strip().splitlines()[-1])
print("%.4g"%d["value"], d["roofline"]["step_frac"], {k:round(v["ms"],2) for k,v in d["roofline"]["kernels"].items()}, d.get("session_pipeline"), d.get("setup_kernel", d.get("setup")), d["clocks"]["sm_mhz"])
PY
