#!/bin/bash
# usage: tools/gpu_prof.sh <tag> <kernel-regex> <n> <count>   (one ncu --set full capture of the 2nd matching launch)
set -u
tag=$1; kern=$2; n=$3; count=$4
mkdir -p gpurun_out
timeout 300 python tools/prof_fold.py $n $count 2 > gpurun_out/prof_$tag.plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/prof_$tag.plain.log; exit 1; }
timeout 900 ncu --set full --import-source on --clock-control none -k regex:$kern -s 1 -c 1 -f -o gpurun_out/prof_$tag \
  python tools/prof_fold.py $n $count 2 > gpurun_out/prof_$tag.ncu.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/prof_$tag.ncu.log; ls -la gpurun_out/prof_$tag.ncu-rep
