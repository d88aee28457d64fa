#!/bin/bash
# GPU box helper: per-kernel durations of the last VMC step of a short bench run (ncu launch list; cold-cache, serialised:
# shares only).  usage: tools/step_launches.sh <workload> <precision>
cmd="python bench.py --workload $1 --precision $2 --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --no-extra"
$cmd > /dev/null 2>&1 || { echo "plain run failed"; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/step_launches.csv $cmd > gpurun_out/step_ncu.log 2>&1
python - <<'PY'
import csv
lines=[l for l in open('gpurun_out/step_launches.csv') if not l.startswith('==')]
rows=[r for r in csv.DictReader(lines) if r['Metric Name']=='gpu__time_duration.sum']
names=[r['Kernel Name'] for r in rows]
# last step: from the last sampling sweep on
last=max(i for i,n in enumerate(names) if 'sweep' in n and i+1 < len(names))
start=max(i for i in range(last) if 'sweep' in names[i] and 'van2' not in names[i]) if any('sweep' in n for n in names[:last]) else 0
for r in rows[start:]:
    print(f"{float(r['Metric Value'].replace(',',''))/1e3:9.1f} us  {r['Kernel Name'][:70]}  grid {r['Grid Size']}")
PY
