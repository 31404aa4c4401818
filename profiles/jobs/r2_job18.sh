#!/bin/bash
# Round 2, GPU job 18: K13 iteration -- parity tests, then timing at 16384^2 (and ncu instruction count when asked).
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_rowstream.py -q -x > gpurun_out/r2_pytest18.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_pytest18.log
timeout 300 python tests/rowstream_probe.py --profile
if [ "$1" = "ncu" ]; then
ncu --metrics smsp__inst_executed.sum,gpu__time_duration.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active,sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed,smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio --clock-control none -k regex:rowstream -s 12 -c 1 python tests/rowstream_probe.py --profile 2>&1 | grep -E "inst_executed|duration|issue_active|pipe_alu|fmaheavy|barrier"
fi
