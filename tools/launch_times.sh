#!/bin/bash
# per-kernel durations of one bench step (ncu launch list), last step only
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_tmp.csv python bench.py --steps 2 --warmup 1 --cpu-pairs 1 --no-em > /dev/null 2>&1
python - <<'PY'
import csv
rows=list(csv.reader(open('gpurun_out/launches_tmp.csv')))
hi=[i for i,r in enumerate(rows) if r and r[0]=="ID"][0]
seen={}
for r in rows[hi+1:]:
    seen[r[4].split('(')[0]]=(r[7],r[8],int(r[-1])/1e6)
for k,v in seen.items(): print("%-28s %-14s %-16s %.3f ms"%(k,v[0],v[1],v[2]))
PY
