#!/bin/bash
# scan path: GPU tests, A/B of the early write-out against the old two-read behaviour, per-kernel split
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -k "scan or parallel_in_time" > gpurun_out/scan_tests.log 2>&1; tail -3 gpurun_out/scan_tests.log
SCAN_TIME_BIG_ONLY=1 python tools/scan_time.py > gpurun_out/scan_new.log 2>&1; cat gpurun_out/scan_new.log
for w in 6 8 10 12 16 20 32; do echo "warps/SM $w"; EKSB_SCAN_WARPS_PER_SM=$w SCAN_TIME_BIG_ONLY=1 python tools/scan_time.py 2>&1 | grep "chunk" | sed -n "2p;4p"; done
echo old; EKSB_SCAN_NO_EARLY=1 EKSB_SCAN_WARPS_PER_SM=32 SCAN_TIME_BIG_ONLY=1 python tools/scan_time.py > gpurun_out/scan_old.log 2>&1; cat gpurun_out/scan_old.log
SCAN_TIME_BIG_ONLY=1 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:scan_ --csv --log-file gpurun_out/scan_launches.csv python tools/scan_time.py > gpurun_out/scan_ncu.log 2>&1
python - <<'PY'
import csv, collections
rows=list(csv.reader(open('gpurun_out/scan_launches.csv')))
i=[k for k,r in enumerate(rows) if r and r[0]=='ID'][0]
h=rows[i]; kn=h.index('Kernel Name'); mn=h.index('Metric Name'); mv=h.index('Metric Value'); idc=h.index('ID')
per=collections.OrderedDict()
for r in rows[i+1:]:
    if len(r)<=mv: continue
    per.setdefault((r[idc], r[kn].split('(')[0][:60]),{})[r[mn]]=float(r[mv].replace(',',''))
for (i_,k),m in list(per.items())[-12:]:
    print(i_, k, {a: round(b/1e6,2) for a,b in m.items()})
PY
