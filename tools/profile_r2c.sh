#!/bin/bash
# ncu captures of the 1D kernels after the narrow-tile / batched-GEMM changes (C1, C1 with 3 layers, C2): launch list + the
# sampling and connected-configuration sweeps + the reverse sweep (same method as tools/profile_r2b.sh)
set -u
OUT=gpurun_out
cap() {
  local tag=$1 wl=$2 kre=$3 skip=$4 cnt=$5
  local cmd="python bench.py --workload $wl --precision auto --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --no-extra"
  if [ ! -f $OUT/${wl}_plain.log ]; then
    $cmd > $OUT/${wl}_plain.log 2>&1 || { echo "$tag: plain run failed"; return; }
    ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/r2c_${wl}_launches.csv $cmd > /dev/null 2>&1
  fi
  ncu --set full --clock-control none --import-source on -k "regex:$kre" -s $skip -c $cnt -f -o $OUT/${tag} $cmd > $OUT/${tag}_ncu2.log 2>&1
  ncu -i $OUT/${tag}.ncu-rep --page raw --csv > $OUT/${tag}_raw.csv 2>/dev/null
  rm -f $OUT/${tag}.ncu-rep $OUT/${tag}_ncu2.log
  echo "== $tag"; python tools/ncu_summary.py $OUT/${tag}_raw.csv
}
cap r2c_c1_sweeps tfim1d_n20_gru50 "^gru1d_sweep_kernel" 2 2
cap r2c_c1_bwd tfim1d_n20_gru50 "^gru1d_bwd_kernel|^gemm_tn_batched_kernel" 2 2
cap r2c_c1x3_sweeps tfim1d_n20_gru50x3 "^gru1d_sweep_kernel" 2 2
cap r2c_c2_sweeps j1j2_n20_cgru50 "^gru1d_sweep_kernel" 2 2
rm -f $OUT/*_plain.log
