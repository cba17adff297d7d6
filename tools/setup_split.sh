#!/bin/bash
mkdir -p gpurun_out
python tools/setup_time.py > gpurun_out/setup_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:setup_\|center_rows --csv --log-file gpurun_out/setup_launches.csv python tools/setup_time.py > gpurun_out/setup_ncu.log 2>&1
cat gpurun_out/setup_plain.log | awk 'NR%3==0'
python - <<'PY'
import csv, collections
rows=list(csv.reader(open('gpurun_out/setup_launches.csv')))
i=[k for k,r in enumerate(rows) if r and r[0]=='ID'][0]
h=rows[i]; kn=h.index('Kernel Name'); mv=h.index('Metric Value')
agg=collections.OrderedDict()
for r in rows[i+1:]:
    if len(r)<=mv: continue
    agg.setdefault(r[kn].split('(')[0][:70],[]).append(float(r[mv].replace(',','')))
for k,v in agg.items(): print(len(v), "first-config per-launch ms:", [round(x/1e6,3) for x in v[:len(v)//9 or 1][:8]], "sum ms per setup call", round(sum(v[:len(v)//9 or 1])/1e6,2), k)
PY
