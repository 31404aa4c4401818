#!/bin/bash
# Round 2, GPU job 14: state check of HEAD at the start of the second session -- whole GPU suite, smoke, default bench (both arms).
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/r2_pytest14.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/r2_pytest14.log
timeout 300 python __graft_entry__.py --smoke > gpurun_out/r2_smoke14.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/r2_smoke14.log
timeout 600 python bench.py > gpurun_out/r2_bench14_default.json 2> gpurun_out/r2_bench14_default.err; echo "bench default rc=$?"
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/r2_bench14_reference.json 2> gpurun_out/r2_bench14_reference.err; echo "bench reference rc=$?"
cat gpurun_out/r2_bench14_default.json | cut -c1-3000
