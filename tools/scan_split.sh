#!/bin/bash
# per-kernel time split of the scan path (ncu launch list; serialised cold-cache times: compare shares)
mkdir -p gpurun_out
python tools/scan_time.py > gpurun_out/scan_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:scan_ --csv --log-file gpurun_out/scan_launches.csv python tools/scan_time.py > gpurun_out/scan_ncu.log 2>&1
grep 10000000 gpurun_out/scan_plain.log
python - <<'PY'
import csv, collections
rows=list(csv.reader(open('gpurun_out/scan_launches.csv')))
i=[k for k,r in enumerate(rows) if r and r[0]=='ID'][0]
h=rows[i]; kn=h.index('Kernel Name'); mv=h.index('Metric Value')
agg=collections.OrderedDict()
for r in rows[i+1:]:
    if len(r)<=mv: continue
    agg.setdefault(r[kn].split('(')[0][:70],[]).append(float(r[mv].replace(',','')))
for k,v in agg.items(): print(len(v), [round(x/1e6,2) for x in v[-3:]], k)
PY
