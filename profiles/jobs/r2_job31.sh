#!/bin/bash
# Round 2, GPU job 31 (2 GPUs): K12 on row slabs (packets to the neighbouring GPU through peer memory): single-GPU slab and
# batch-resident tests first, then the multi-GPU slab check and the slab bench; replica-exchange parity for the epoch change.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_slab.py tests/test_zx_gpu_pbatch.py tests/test_gpu_rx.py tests/test_zz_no_alloc.py -q -x > gpurun_out/r2_pytest31.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2_pytest31.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
timeout 300 $TR tests/mgpu_slab_check.py --bench > gpurun_out/r2_mgpu_slab31.log 2>&1; echo "mgpu slab rc=$?"; grep -v "^\*\|OMP_NUM" gpurun_out/r2_mgpu_slab31.log | tail -12
timeout 300 $TR tests/mgpu_rx_check.py > gpurun_out/r2_mgpu_rx31.log 2>&1; echo "mgpu rx rc=$?"; tail -2 gpurun_out/r2_mgpu_rx31.log
timeout 300 $TR bench.py --gpus 2 --workload slab --steps 200 --warmup 20 > gpurun_out/r2_bench31_n2_slab.json 2> gpurun_out/r2_bench31_n2_slab.err; echo "bench slab rc=$?"; tail -2 gpurun_out/r2_bench31_n2_slab.err; cut -c1-400 gpurun_out/r2_bench31_n2_slab.json
