#!/bin/bash
# Round 2, GPU job 1: full GPU test-suite with the un-gated persistent-kernel cases, smoke, size/kT probe of the
# persistent kernel, default bench line, launch list and the ncu --set full capture of persistent_kernel.
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/r2_pytest1.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/r2_pytest1.log
python __graft_entry__.py --smoke > gpurun_out/r2_smoke1.log 2>&1; echo "smoke rc=$?"
python tests/persist_probe.py --out gpurun_out/r2_persist_probe.json > gpurun_out/r2_probe.log 2>&1; echo "probe rc=$?"
python bench.py --steps 20 --warmup 3 > gpurun_out/r2_bench1.json 2> gpurun_out/r2_bench1.err; echo "bench rc=$?"
CMD="python bench.py --steps 20 --warmup 3 --no-cpu --no-also-bits --e2e-sweeps 4"
$CMD > gpurun_out/r2_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches.csv $CMD > gpurun_out/r2_ncu_launch.log 2>&1
echo "launch list rc=$?"
$CMD > gpurun_out/r2_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:persistent -c 2 -o gpurun_out/r2_persist $CMD > gpurun_out/r2_ncu_full.log 2>&1
echo "ncu full rc=$?"
