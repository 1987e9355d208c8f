#!/bin/bash
for so in build_var/clk.so build_var/skip*.so; do
  echo "== $so"
  RNAFOLD_LIB=$PWD/$so timeout 120 python tools/phase_clocks.py ${1:-100} ${2:-23680} 2>&1 | awk 'NR==1 || /d  20|d  44|d  76|d  92/'
done
