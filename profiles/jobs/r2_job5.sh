#!/bin/bash
# Round 2, GPU job 5: ncu --set full of the batch-resident kernel on the bench headline (16384^2 int8, 20 sweeps per
# launch) and on the replica-exchange batch (256 x 1024^2, exchange every 10 sweeps inside the kernel).
mkdir -p gpurun_out
CMD="python bench.py --steps 20 --warmup 5 --no-cpu --no-extras --e2e-sweeps 4"
$CMD > gpurun_out/r2_plain5.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:pbatch -s 2 -c 1 -o gpurun_out/r2_pbatch $CMD > gpurun_out/r2_ncu5.log 2>&1
echo "ncu single rc=$?"
CMD2="python bench.py --workload rx --rx-replicas-total 256 --steps 20 --warmup 10 --no-cpu --no-extras --e2e-sweeps 10"
$CMD2 > gpurun_out/r2_plain5b.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:pbatch -s 2 -c 1 -o gpurun_out/r2_pbatch_rx $CMD2 > gpurun_out/r2_ncu5b.log 2>&1
echo "ncu rx rc=$?"
cat gpurun_out/r2_plain5b.log | tail -2
