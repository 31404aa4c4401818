#!/bin/bash
# Round 2, GPU job 30: final build -- whole GPU suite, smoke, the driver's bench command (both arms), launch list, ncu --set full of K13.
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/r2_pytest30.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_pytest30.log
timeout 300 python __graft_entry__.py --smoke > gpurun_out/r2_smoke30.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/r2_smoke30.log
timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2_bench30_default.json 2> gpurun_out/r2_bench30_default.err; echo "bench default rc=$?"
timeout 600 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2_bench30_reference.json 2> gpurun_out/r2_bench30_reference.err; echo "bench reference rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2_bench30_default.json').read())
print('value %.4e'%d['value'], 'frac', d['roofline']['frac'], 'e2e %.4e'%d['e2e']['value'], 'serial %.3e'%d['e2e']['serial']['value'])
print('one', d['one_sweep_per_launch']['ms_per_step'], d['one_sweep_per_launch']['roofline_frac'], d['one_sweep_per_launch']['kernel'], 'bits %.4e'%d['bitpacked']['value'], 'rx %.4e'%d['rx']['value'], d['rx']['exchange_us'])
r=json.loads(open('gpurun_out/r2_bench30_reference.json').read()); print('reference %.4e'%r['value'], r['cpu_baseline']['kind'])
PY
CMD="python bench.py --steps 20 --warmup 5 --no-cpu --no-also-bits --e2e-sweeps 4 --e2e-depth 2"
$CMD > gpurun_out/r2_plain30.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches30.csv $CMD > gpurun_out/r2_ncu_launch30.log 2>&1; echo "launch list rc=$?"
CMD="python tests/rowstream_probe.py --profile"
$CMD > gpurun_out/r2_plain30b.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:rowstream -s 12 -c 1 -o gpurun_out/r2_rowstream30 $CMD > gpurun_out/r2_ncu30b.log 2>&1; echo "ncu rowstream rc=$?"
