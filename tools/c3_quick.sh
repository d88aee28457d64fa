#!/bin/bash
# GPU box helper: vanilla split-precision parity tests + three short bench runs of the C3 configuration
timeout 200 python -m pytest tests/test_gpu_tc3.py tests/test_gpu_bench_size.py -q -m gpu -k "vanilla or tfim2d or van" 2>&1 | tail -2
for rep in 1 2 3; do
  timeout 120 python bench.py --workload tfim2d_12x12_rnn50 --precision tc3 --steps 10 --warmup 3 --no-cpu-baseline --no-e2e --no-extra 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(d['precision_mode'], round(d['ms_per_step'],2), {k: round(v,2) for k,v in d['phase_ms_per_step'].items()}, round(d['roofline']['frac'],3), d['clocks']['sm_mhz'])"
done
