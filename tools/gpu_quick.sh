#!/bin/bash
# Short GPU session: parity tests, kernel timing probe, one bench line per dtype.
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee gpurun_out/rc.txt
timeout 600 python tools/quick_time.py > gpurun_out/quick_time.log 2>&1; echo "quick rc=$?" | tee -a gpurun_out/rc.txt
timeout 600 python bench.py --no-cpu --no-e2e > gpurun_out/bench_f64_k.json 2> gpurun_out/bench_f64_k.err; echo "bench f64 rc=$?" | tee -a gpurun_out/rc.txt
timeout 600 python bench.py --dtype f32 --no-cpu --no-e2e > gpurun_out/bench_f32_k.json 2> gpurun_out/bench_f32_k.err; echo "bench f32 rc=$?" | tee -a gpurun_out/rc.txt
tail -3 gpurun_out/pytest_gpu.log; cat gpurun_out/quick_time.log
python - <<'PY'
import json
for f in ("bench_f64_k","bench_f32_k"):
    try:
        d=json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
        print(f, d["value"], d["roofline"]["step_frac"], {k:(round(v["ms"],2),round(v["frac"],3)) for k,v in d["roofline"]["kernels"].items()})
    except Exception as e: print(f, "ERR", e)
PY
