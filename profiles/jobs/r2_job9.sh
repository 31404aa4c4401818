#!/bin/bash
# Round 2, GPU job 9 (2 GPUs): bench.py under torchrun at N = 2 (config 3 strong + weak + in-job N = 1 check), the
# multi-GPU parity scripts, and the reference arm launched the same way.
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r2_bench9_n2.json 2> gpurun_out/r2_bench9_n2.err; echo "bench n2 rc=$?"
tail -5 gpurun_out/r2_bench9_n2.err
cat gpurun_out/r2_bench9_n2.json
timeout 600 $TR tests/mgpu_rx_check.py > gpurun_out/r2_mgpu_rx9.log 2>&1; echo "mgpu rx rc=$?"; tail -3 gpurun_out/r2_mgpu_rx9.log
timeout 600 $TR tests/mgpu_slab_check.py > gpurun_out/r2_mgpu_slab9.log 2>&1; echo "mgpu slab rc=$?"; tail -3 gpurun_out/r2_mgpu_slab9.log
