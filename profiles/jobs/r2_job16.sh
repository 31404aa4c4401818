#!/bin/bash
# Round 2, GPU job 16: first run of the row-streaming kernel K13 -- parity tests (short timeout: a wrong mbarrier would
# spin until the kernel's own two-second bound), then the probe against the tiled kernel.
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_rowstream.py -q -x > gpurun_out/r2_pytest16.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/r2_pytest16.log
timeout 600 python tests/rowstream_probe.py --out gpurun_out/r2_rowstream_probe16.json > gpurun_out/r2_probe16.log 2>&1; echo "probe rc=$?"; tail -30 gpurun_out/r2_probe16.log
