#!/bin/bash
# Per-kernel durations of one step (ncu launch list; cold-cache, serialised): tools/launch_times.sh [out.csv]
OUT=${1:-gpurun_out/launches.csv}
timeout 500 ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none -s 32 -c 16 --csv --log-file $OUT python tools/step_breakdown.py 1024 328 2 > gpurun_out/ncu_l.log 2>&1
python - "$OUT" <<'PY'
import csv, sys
rows=[r for r in csv.reader(open(sys.argv[1])) if len(r)>10]
h=rows[0]; i=h.index("Kernel Name"); m=h.index("Metric Name"); v=h.index("Metric Value")
cur={}
for r in rows[1:]:
    k=r[i].split('(')[0][-28:]
    if r[m].startswith('gpu__time'): t=float(r[v])/1e3
    else: print(f"{k:30s} {t:9.1f} us {float(r[v])/1e6:9.1f} Minst")
PY
