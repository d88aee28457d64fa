#!/bin/bash
# GPU box helper: split-precision parity tests, then old / new library on the C4 and C5 configurations in tc3
timeout 400 python -m pytest tests/test_gpu_tc3.py tests/test_gpu_bench_size.py -q -m gpu 2>&1 | tail -2
for wl in heis10x10_tri_tgru100 rydberg_4x4_gru256; do for lib in old new; do
  if [ $lib = old ]; then export RNNWF_LIB_PATH=$PWD/rnnwavefunctions_b200/librnnwf_old.so; else unset RNNWF_LIB_PATH; fi
  timeout 200 python bench.py --workload $wl --precision tc3 --steps 5 --warmup 3 --no-cpu-baseline --no-e2e --no-extra 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$wl $lib', d['precision_mode'], round(d['ms_per_step'],2), {k: round(v,2) for k,v in d['phase_ms_per_step'].items()}, round(d['roofline']['frac'],3), d['clocks']['sm_mhz'])"
done; done
