#!/bin/bash
# A/B of ring shapes in the bench's sustained (power-capped) regime
run() { echo "== $*"; env "$@" python bench.py --no-cpu --no-e2e --no-setup --no-sustained --steps 20 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d['roofline']
print('%.4g'%d['value'], round(r['step_frac'],3), {k:round(v['ms'],2) for k,v in r['kernels'].items()}, d['clocks']['sm_mhz'], d['clocks']['power_w_max'])"; }
run A=1
run EKSB_FWD_SHALLOW=1
run EKSB_BWD_NARROW=1
run EKSB_FWD3_MAX=2
run A=1
