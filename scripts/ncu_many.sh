#!/bin/bash
# ncu --set full with source for a list of "tag kernel-regex kind H Cin N taps"; CSV exports land in gpurun_out/
while read tag rx kind h c n t; do
  [ -z "$tag" ] && continue
  python scripts/gemm_one.py $kind $h $c $n $t > /dev/null 2>&1 || { echo "$tag plain run failed"; continue; }
  ncu --set full --clock-control none --import-source on -k regex:$rx -s 2 -c 1 -o /tmp/p_$tag -f python scripts/gemm_one.py $kind $h $c $n $t > gpurun_out/ncu_$tag.log 2>&1
  ncu -i /tmp/p_$tag.ncu-rep --page raw --csv > gpurun_out/prof_${tag}_raw.csv
  ncu -i /tmp/p_$tag.ncu-rep --page source --csv > gpurun_out/prof_${tag}_src.csv 2>/dev/null
  echo "$tag done"
done
