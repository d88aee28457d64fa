#!/bin/bash
# A/B of two builds of the library on the same box: tools/ab.sh "<bench args>"  (old = rnnwavefunctions_b200/librnnwf_old.so)
for rep in 1 2; do for lib in old new; do
  if [ $lib = old ]; then export RNNWF_LIB_PATH=$PWD/rnnwavefunctions_b200/librnnwf_old.so; else unset RNNWF_LIB_PATH; fi
  python bench.py $1 --steps 10 --warmup 3 --no-cpu-baseline --no-e2e --no-extra 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$lib', d['precision_mode'], round(d['ms_per_step'],2), {k: round(v,2) for k,v in d['phase_ms_per_step'].items()}, round(d['roofline']['frac'],3), d['clocks']['sm_mhz'])"
done; done
