#!/bin/bash
# instruction counts + duration of the wave kernel for each variant library (ncu, metrics only)
for so in rnagent_b200/librnafold_b200.so build_var/*.so; do
  [ -f "$so" ] || continue
  RNAFOLD_LIB=$PWD/$so timeout 300 ncu --metrics smsp__inst_executed.sum,gpu__time_duration.sum,launch__registers_per_thread,smsp__issue_active.avg.pct_of_peak_sustained_active,smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio,smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio,smsp__average_warps_issue_stalled_wait_per_issue_active.ratio,smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio --clock-control none -k regex:fold_wave -s 1 -c 1 --csv python tools/prof_fold.py ${VAR_N:-100} ${VAR_COUNT:-23680} 2 2>&1 | grep -E "inst_executed|time_duration|registers|issue_active|no_instruction|stalled_barrier|stalled_wait|stalled_short" | awk -F'","' -v so=$so '{print so, $(NF-2), $NF}'
done
