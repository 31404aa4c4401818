#!/bin/bash
# Round 2, GPU job 27: bench lines of the final build (the driver's command, both arms; configs 3-5 on one GPU; the shipped
# ladder with an exchange every sweep), launch list and ncu --set full captures of K12 (single lattice and replica-exchange
# batch) and K13 with that build.
mkdir -p gpurun_out
timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2_bench27_default.json 2> gpurun_out/r2_bench27_default.err; echo "bench default rc=$?"
timeout 600 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2_bench27_reference.json 2> gpurun_out/r2_bench27_reference.err; echo "bench reference rc=$?"
for W in "latgas" "latgas --kernel two_pass" "discs" "rx --rx-model hardcore --rx-replicas-total 64 --rx-kT-lo 0.5 --rx-kT-hi 2.0" "rx --rx-kT-lo 0.01 --rx-kT-hi 5.0 --rx-every 1"; do
  tag=$(echo $W | tr ' -.' '___')
  timeout 600 python bench.py --workload $W --steps 20 --warmup 5 > gpurun_out/r2_bench27_$tag.json 2> gpurun_out/r2_bench27_$tag.err; echo "bench $W rc=$?"
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_bench27_*.json')):
    try:
        d=json.loads([l for l in open(f).read().splitlines() if l.startswith('{')][-1])
        print(f.split('r2_bench27_')[1], '%.4e'%d['value'], 'ms/step', round(d['ms_per_step'],5), 'frac', d.get('roofline',{}).get('frac'), 'e2e', d.get('e2e',{}).get('value'), d.get('parity'), d.get('exchange_us'))
    except Exception as e: print(f, 'ERR', e)
d=json.loads(open('gpurun_out/r2_bench27_default.json').read())
print('one', d['one_sweep_per_launch']['ms_per_step'], d['one_sweep_per_launch']['roofline_frac'], 'bits', d['bitpacked']['value'], 'rx', d['rx']['value'], d['rx']['exchange_us'])
PY
CMD="python bench.py --steps 20 --warmup 5 --no-cpu --no-also-bits --e2e-sweeps 4 --e2e-depth 2"
$CMD > gpurun_out/r2_plain27.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches27.csv $CMD > gpurun_out/r2_ncu_launch27.log 2>&1; echo "launch list rc=$?"
CMD="python bench.py --steps 20 --warmup 5 --no-cpu --no-extras --e2e-sweeps 4 --e2e-depth 2"
$CMD > gpurun_out/r2_plain27b.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:pbatch -s 2 -c 1 -o gpurun_out/r2_pbatch27 $CMD > gpurun_out/r2_ncu27b.log 2>&1; echo "ncu single rc=$?"
CMD2="python bench.py --workload rx --rx-replicas-total 256 --steps 20 --warmup 10 --no-cpu --no-extras --e2e-sweeps 10"
$CMD2 > gpurun_out/r2_plain27c.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:pbatch -s 6 -c 1 -o gpurun_out/r2_pbatch_rx27 $CMD2 > gpurun_out/r2_ncu27c.log 2>&1; echo "ncu rx rc=$?"
