#!/bin/bash
# per-kernel average durations (ncu, serialised): usage tools/klist.sh TAG [env...]
T=$1; shift
env "$@" ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${T}_kl.csv python tools/prof_run.py --bands ${KL_BANDS:-32} --calls 2 $KL_ARGS > /dev/null 2>&1
python - <<PY
import csv,collections
rows=[r for r in csv.reader(open('gpurun_out/${T}_kl.csv')) if len(r)>10 and r[0].isdigit()]
agg=collections.OrderedDict()
for r in rows:
    k=r[4].split('(')[0][:60]
    if 'at::' in k: continue
    a=agg.setdefault(k,[0,0.0,r[7],r[8]]); a[0]+=1; a[1]+=float(r[-1])
print('== ${T}', '$*')
for k,a in agg.items(): print(f"{k:60s} n={a[0]:3d} avg={a[1]/a[0]/1e3:8.2f}us blk={a[2]} grid={a[3]}")
PY
