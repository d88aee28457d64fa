#!/bin/bash
# GPU box helper: per-kernel durations of one C4 backward (ncu launch list; cold-cache, serialised: shares only)
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/bwd_launches.csv python tools/bwd_timing.py 200 > gpurun_out/bwd_ncu.log 2>&1
python - <<'PY'
import csv
lines=[l for l in open('gpurun_out/bwd_launches.csv') if not l.startswith('==')]
rows=[r for r in csv.DictReader(lines) if r['Metric Name']=='gpu__time_duration.sum']
for r in rows[-40:]:
    print(f"{float(r['Metric Value'].replace(',',''))/1e3:9.1f} us  {r['Kernel Name'][:60]}  grid {r['Grid Size']}")
PY
