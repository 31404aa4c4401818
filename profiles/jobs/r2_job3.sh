#!/bin/bash
# Round 2, GPU job 3: K12 after the wait / publish rewrite (acquire loads, back-off, abort word in device memory):
# its tests, the timing probe, and the rewritten bench.py at N = 1 (both arms).
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_zx_gpu_pbatch.py -q > gpurun_out/r2_pytest_pbatch3.log 2>&1; echo "pbatch tests rc=$?"
tail -5 gpurun_out/r2_pytest_pbatch3.log
timeout 600 python tests/pbatch_probe.py --out gpurun_out/r2_pbatch_probe3.json > gpurun_out/r2_pbatch_probe3.log 2>&1; echo "probe rc=$?"
tail -12 gpurun_out/r2_pbatch_probe3.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r2_bench3.json 2> gpurun_out/r2_bench3.err; echo "bench rc=$?"
tail -3 gpurun_out/r2_bench3.err
cat gpurun_out/r2_bench3.json
timeout 600 python bench.py --impl reference --steps 4 --warmup 1 > gpurun_out/r2_bench3_ref.json 2> gpurun_out/r2_bench3_ref.err; echo "ref rc=$?"
