#!/bin/bash
# Round 2, GPU job 40: final build on one GPU -- whole GPU suite, smoke, the driver's bench command (both arms), launch list, ncu --set
# full captures of K12 (single lattice, replica-exchange batch); configs 4 / 5 and the shipped ladder on one GPU.
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/r2_pytest40.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_pytest40.log
timeout 300 python __graft_entry__.py --smoke > gpurun_out/r2_smoke40.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/r2_smoke40.log
timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2_bench40_default.json 2> gpurun_out/r2_bench40_default.err; echo "bench default rc=$?"
timeout 600 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2_bench40_reference.json 2> gpurun_out/r2_bench40_reference.err; echo "bench reference rc=$?"
for W in "latgas" "discs" "rx --rx-model hardcore --rx-replicas-total 64 --rx-kT-lo 0.5 --rx-kT-hi 2.0" "rx --rx-kT-lo 0.01 --rx-kT-hi 5.0 --rx-every 1"; do
  tag=$(echo $W | tr ' -.' '___')
  timeout 600 python bench.py --workload $W --steps 20 --warmup 5 > gpurun_out/r2_bench40_$tag.json 2> gpurun_out/r2_bench40_$tag.err; echo "bench $W rc=$?"
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_bench40_*.json')):
    try:
        d=json.loads([l for l in open(f).read().splitlines() if l.startswith('{')][-1])
        print(f.split('r2_bench40_')[1], '%.4e'%d['value'], 'ms/step', round(d['ms_per_step'],5), 'frac', d.get('roofline',{}).get('frac'), 'e2e', d.get('e2e',{}).get('value'), d.get('parity'), d.get('exchange_us'))
    except Exception as e: print(f, 'ERR', e)
d=json.loads(open('gpurun_out/r2_bench40_default.json').read())
print('one', d['one_sweep_per_launch']['ms_per_step'], d['one_sweep_per_launch']['roofline_frac'], 'bits %.4e'%d['bitpacked']['value'], 'rx %.4e'%d['rx']['value'], d['rx']['exchange_us'], 'serial %.3e'%d['e2e']['serial']['value'])
PY
CMD="python bench.py --steps 20 --warmup 5 --no-cpu --no-also-bits --e2e-sweeps 4 --e2e-depth 2"
$CMD > gpurun_out/r2_plain40.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches40.csv $CMD > gpurun_out/r2_ncu_launch40.log 2>&1; echo "launch list rc=$?"
CMD="python bench.py --steps 20 --warmup 5 --no-cpu --no-extras --e2e-sweeps 4 --e2e-depth 2"
$CMD > gpurun_out/r2_plain40b.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:pbatch -s 2 -c 1 -o gpurun_out/r2_pbatch40 $CMD > gpurun_out/r2_ncu40b.log 2>&1; echo "ncu single rc=$?"
CMD2="python bench.py --workload rx --rx-replicas-total 256 --steps 20 --warmup 10 --no-cpu --no-extras --e2e-sweeps 10"
$CMD2 > gpurun_out/r2_plain40c.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:pbatch -s 6 -c 1 -o gpurun_out/r2_pbatch_rx40 $CMD2 > gpurun_out/r2_ncu40c.log 2>&1; echo "ncu rx rc=$?"
