#!/bin/bash
# Round 2, GPU job 45 (one GPU): K12 with two blocks per SM where both fit (PYMC_PBATCH_BLOCKS_PER_SM=2) against one;
# parity files of K12 and the slabs with two.
mkdir -p gpurun_out
for P in 1 2; do
  export PYMC_PBATCH_BLOCKS_PER_SM=$P
  echo "== blocks per SM $P"
  for R in 8 32 64; do timeout 120 python tests/k12_small_probe.py $R; done
  for L in 128 1024 2048 4096; do timeout 120 python tests/k12_single_probe.py $L; done
done 2>&1 | tee gpurun_out/r2_pair45.log
export PYMC_PBATCH_BLOCKS_PER_SM=2
timeout 600 python -m pytest tests/test_zx_gpu_pbatch.py tests/test_gpu_slab.py tests/test_gpu_rx.py tests/test_gpu_errors.py -m gpu -x -q 2>&1 | tail -5 | tee gpurun_out/r2_pytest45.log
