#!/bin/bash
# Round 2, GPU job 19: A/B of K13 build variants (libraries under py_mc_b200/_lib/variants/, loaded through PYMC_B200_LIB):
# timing, instruction count and stall reasons per variant.
mkdir -p gpurun_out
M="smsp__inst_executed.sum,gpu__time_duration.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active,sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed,sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active,smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio,smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio,smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio,smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio,smsp__average_warps_issue_stalled_wait_per_issue_active.ratio,smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio,smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio,smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio,smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio,l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"
for v in "$@"; do
  export PYMC_B200_LIB=$PWD/py_mc_b200/_lib/variants/lib_$v.so
  echo "== $v"
  timeout 300 python tests/rowstream_probe.py --profile
  ncu --metrics $M --clock-control none -k regex:rowstream -s 12 -c 1 python tests/rowstream_probe.py --profile 2>&1 | grep -E "^\s+(smsp|sm__|gpu__|l1tex)" | sed 's/smsp__average_warps_issue_stalled_//; s/_per_issue_active.ratio//' | awk '{printf "%s=%s  ", $1, $NF} END{print ""}'
done
