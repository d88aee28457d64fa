#!/bin/bash
# GPU box helper: the whole GPU suite, then the three small-model bench lines (phases, FFMA fraction, launches per step)
python -m pytest tests -q -m gpu > gpurun_out/full.log 2>&1; tail -3 gpurun_out/full.log; grep -E "^FAILED|^ERROR" gpurun_out/full.log | head
for w in tfim1d_n20_gru50 tfim1d_n20_gru50x3 j1j2_n20_cgru50; do
  python bench.py --workload $w --steps 20 --warmup 3 --no-cpu-baseline --no-e2e 2>&1 | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(d['config']['workload'][:24], d['ms_per_step'], d['phase_ms_per_step'], d['roofline']['frac'], d['gpu_launches']/20)"
done
