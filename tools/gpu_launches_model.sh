#!/bin/bash
# ncu launch list of another config: TKB_MODEL / TKB_BATCH
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
TKB_PLAN=1 python tools/profile_step.py tc 0 3 > gpurun_out/plain_m.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_m.csv python tools/profile_step.py tc 0 3 > gpurun_out/ncu_m.log 2>&1
tail -n 2 gpurun_out/ncu_m.log
python - <<'PY'
import csv, collections
lines=[l for l in open('gpurun_out/launches_m.csv') if not l.startswith('==')]
rows=list(csv.DictReader(lines)); n=len(rows)//3
agg=collections.defaultdict(lambda:[0,0.0]); tot=0
for r in rows[2*n:]:
    t=float(r['Metric Value'].replace(',','')); u=r['Metric Unit']; t=t/1000 if u=='ns' else (t*1000 if u=='ms' else t)
    k=r['Kernel Name'][:60]; agg[k][0]+=1; agg[k][1]+=t; tot+=t
for k,(c,t) in sorted(agg.items(), key=lambda kv:-kv[1][1])[:14]: print('%8.1f us %4d x %s'%(t,c,k))
print('total', round(tot,1))
PY
