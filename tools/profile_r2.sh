#!/bin/bash
# Round-2 ncu captures (run on the GPU box through gpurun): for every BASELINE configuration the launch list of a short bench
# run and one `--set full` capture of its connected-configuration sweep (the dominant kernel), exported as raw CSV.
# Each ncu run only follows a plain run of the same command that exited 0.
set -u
OUT=gpurun_out
prof() {   # tag workload precision kernel-regex
  local tag=$1 wl=$2 prec=$3 kre=$4
  local cmd="python bench.py --workload $wl --precision $prec --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --no-extra"
  $cmd > $OUT/${tag}_plain.log 2>&1 || { echo "$tag: plain run failed"; return; }
  ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/${tag}_launches.csv $cmd > $OUT/${tag}_ncu1.log 2>&1
  ncu --set full --clock-control none --import-source on -k regex:$kre -s 2 -c 2 -f -o $OUT/${tag} $cmd > $OUT/${tag}_ncu2.log 2>&1
  ncu -i $OUT/${tag}.ncu-rep --page raw --csv > $OUT/${tag}_raw.csv 2>/dev/null
  ncu -i $OUT/${tag}.ncu-rep --page details --csv 2>/dev/null | grep -E "Stall|stall|Issue|Eligible|Executed Ipc|Theoretical Occ|Achieved Occ|Registers|Shared Memory Config|Duration|SM Frequency|Warp Cycles|No Eligible" > $OUT/${tag}_details.csv
  rm -f $OUT/${tag}.ncu-rep $OUT/${tag}_ncu1.log $OUT/${tag}_ncu2.log      # (the reports with imported source are ~15 MB each)
  echo "$tag done: $(tail -1 $OUT/${tag}_plain.log | cut -c1-120)"
}
prof r2_c1 tfim1d_n20_gru50 auto gru1d_sweep_kernel
prof r2_c1x3 tfim1d_n20_gru50x3 auto gru1d_sweep_kernel
prof r2_c2 j1j2_n20_cgru50 auto gru1d_sweep_kernel
prof r2_c3 tfim2d_12x12_rnn50 tc3 sweep_tc_kernel
prof r2_c4tc heis10x10_tri_tgru100 tc sweep_tc_kernel
prof r2_c4tc3 heis10x10_tri_tgru100 tc3 sweep_tc_kernel
