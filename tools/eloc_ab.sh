#!/bin/bash
# GPU box helper: A/B of a sweep_tc_kernel variant switched by RNNWF_TC_DEBUG=$1 on the headline workload (a few repetitions)
for rep in 1 2 3; do for dbg in 0 $1; do
  RNNWF_TC_DEBUG=$dbg timeout 120 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e --no-extra 2>&1 | tail -1 > /tmp/l.json
  echo -n "debug=$dbg  "; python tools/show_line.py /tmp/l.json | head -1
done; done
