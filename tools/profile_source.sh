#!/bin/bash
# GPU box helper: source-level ncu capture of one kernel launch -> gpurun_out/<tag>_source.csv
# usage: tools/profile_source.sh <tag> <workload> <precision> <kernel-regex> <skip>
set -u
OUT=gpurun_out; tag=$1; wl=$2; prec=$3; kre=$4; skip=$5
cmd="python bench.py --workload $wl --precision $prec --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --no-extra"
$cmd > /dev/null 2>&1 || { echo "plain run failed"; exit 1; }
ncu --set full --section SourceCounters --clock-control none --import-source on -k "regex:$kre" -s $skip -c 1 -f -o $OUT/$tag $cmd > $OUT/${tag}_ncu.log 2>&1
ncu -i $OUT/$tag.ncu-rep --page source --csv > $OUT/${tag}_source.csv 2>/dev/null
ncu -i $OUT/$tag.ncu-rep --page raw --csv > $OUT/${tag}_raw.csv 2>/dev/null
rm -f $OUT/$tag.ncu-rep
wc -c $OUT/${tag}_source.csv
