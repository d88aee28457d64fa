#!/bin/bash
# Round-2 (second half) ncu captures, run on the GPU box through gpurun: the kernels changed after profiles/r2_summary.md --
# C4 backward (reverse sweep, weight gradients), C4 sampling sweep, C3 tc3 sweep, C5 sweeps (tc and tc3) -- plus a fresh
# capture of the headline sweep.  Each ncu run follows a plain run of the same command that exited 0.
set -u
OUT=gpurun_out
cap() {   # tag workload precision kernel-regex skip count   (sweeps: a step launches the sampling instance, then the CHILD one)
  local tag=$1 wl=$2 prec=$3 kre=$4 skip=$5 cnt=$6
  local cmd="python bench.py --workload $wl --precision $prec --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --no-extra"
  if [ ! -f $OUT/${wl}_${prec}_plain.log ]; then
    $cmd > $OUT/${wl}_${prec}_plain.log 2>&1 || { echo "$tag: plain run failed"; return; }
    ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/r2b_${wl}_${prec}_launches.csv $cmd > /dev/null 2>&1
  fi
  ncu --set full --clock-control none --import-source on -k "regex:$kre" -s $skip -c $cnt -f -o $OUT/${tag} $cmd > $OUT/${tag}_ncu2.log 2>&1
  ncu -i $OUT/${tag}.ncu-rep --page raw --csv > $OUT/${tag}_raw.csv 2>/dev/null
  ncu -i $OUT/${tag}.ncu-rep --page details --csv 2>/dev/null | grep -E "Stall|stall|Issue|Eligible|Executed Ipc|Theoretical Occ|Achieved Occ|Registers|Duration|SM Frequency|Warp Cycles|No Eligible" > $OUT/${tag}_details.csv
  rm -f $OUT/${tag}.ncu-rep $OUT/${tag}_ncu2.log
  echo "== $tag"; python tools/ncu_summary.py $OUT/${tag}_raw.csv
}
cap r2b_c4_sweeps heis10x10_tri_tgru100 tc "^sweep_tc_kernel" 2 2
cap r2b_c4_bwd heis10x10_tri_tgru100 tc "^bwd_tc_kernel" 1 1
cap r2b_c3_sweeps tfim2d_12x12_rnn50 tc3 "^sweep_tc_kernel" 2 2
cap r2b_c5_sweeps rydberg_4x4_gru256 tc "^sweep_tc4_kernel" 2 2
cap r2b_c5tc3_sweeps rydberg_4x4_gru256 tc3 "^sweep_tc4_kernel" 2 2
rm -f $OUT/*_plain.log
