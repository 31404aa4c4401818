#!/bin/bash
# Round 2, GPU job 17: ncu --set full of the first version of the row-streaming kernel K13 (16384^2 int8, one sweep per launch).
mkdir -p gpurun_out
CMD="python tests/rowstream_probe.py --profile"
$CMD > gpurun_out/r2_plain17.log 2>&1 && cat gpurun_out/r2_plain17.log &&
ncu --set full --clock-control none --import-source on -k regex:rowstream -s 12 -c 1 -o gpurun_out/r2_rowstream17 $CMD > gpurun_out/r2_ncu17.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/r2_ncu17.log
