#!/bin/bash
# Round 2, GPU job 2: first run of the batch-resident kernel K12 -- its own tests first (bounded), then the whole
# GPU suite, smoke, and the timing probe.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_zx_gpu_pbatch.py -q -x > gpurun_out/r2_pytest_pbatch.log 2>&1; echo "pbatch tests rc=$?"
tail -15 gpurun_out/r2_pytest_pbatch.log
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/r2_pytest2.log 2>&1; echo "pytest rc=$?"
tail -8 gpurun_out/r2_pytest2.log
timeout 300 python __graft_entry__.py --smoke > gpurun_out/r2_smoke2.log 2>&1; echo "smoke rc=$?"
timeout 600 python tests/pbatch_probe.py --out gpurun_out/r2_pbatch_probe.json > gpurun_out/r2_pbatch_probe.log 2>&1; echo "probe rc=$?"
tail -12 gpurun_out/r2_pbatch_probe.log
