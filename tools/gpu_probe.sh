#!/bin/bash
# sustained-run clock/power trace + per-kernel time split of the scan path
mkdir -p gpurun_out
nvidia-smi --query-gpu=power.limit,power.max_limit,power.default_limit,clocks.max.sm,clocks.max.mem --format=csv
(nvidia-smi --query-gpu=clocks.sm,clocks.mem,power.draw,temperature.gpu,clocks_event_reasons.sw_power_cap,clocks_event_reasons.sw_thermal_slowdown --format=csv,noheader -lms 100 > gpurun_out/smi_trace.csv &)
REPS=12 python tools/quick_time.py 64:8192:100000
sleep 0.3
awk "NR%4==0" gpurun_out/smi_trace.csv | head -40
python tools/scan_time.py > gpurun_out/scan_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:scan_ --csv --log-file gpurun_out/scan_launches.csv python tools/scan_time.py > gpurun_out/scan_ncu.log 2>&1
python - <<'PY'
import csv, collections
rows=list(csv.reader(open('gpurun_out/scan_launches.csv')))
i=[k for k,r in enumerate(rows) if r and r[0]=='ID'][0]
h=rows[i]; kn=h.index('Kernel Name'); mv=h.index('Metric Value'); 
agg=collections.OrderedDict()
for r in rows[i+1:]:
    if len(r)<=mv: continue
    name=r[kn].split('(')[0][:70]
    agg.setdefault(name,[]).append(float(r[mv].replace(',','')))
for k,v in agg.items(): print(len(v), [round(x/1e6,2) for x in v[-4:]], k)
PY
