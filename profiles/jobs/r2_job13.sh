#!/bin/bash
# Round 2, GPU job 13: K12 with the lean path for all warps without export items; GPU suite, probe, bench, and the ncu
# capture of K12 on the replica-exchange batch (256 x 1024^2, 20 sweeps + 2 exchanges per launch).
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/r2_pytest13.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2_pytest13.log
timeout 600 python tests/pbatch_probe.py --out gpurun_out/r2_pbatch_probe13.json > gpurun_out/r2_pbatch_probe13.log 2>&1; echo "probe rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2_pbatch_probe13.json'))
for L,row in d['single'].items(): print(L, {k:round(v.get('ms_per_sweep',-1)*1e3,2) for k,v in row.items() if 'pbatch' in k or 'persistent' in k})
for R,row in d['rx'].items(): print(R, {k:round(v.get('ms_per_sweep',-1)*1e3,2) for k,v in row.items() if 'pbatch' in k})
print(d['latgas'])
PY
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu > gpurun_out/r2_bench13.json 2> gpurun_out/r2_bench13.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r2_bench13.json') if l.startswith('{')][-1])
print(d['value'], d['ms_per_step'], d['rx']['value'], d['rx']['exchange_us'], d['bitpacked']['value'])
PY
CMD2="python bench.py --workload rx --rx-replicas-total 256 --steps 20 --warmup 10 --no-cpu --no-extras --e2e-sweeps 10"
$CMD2 > gpurun_out/r2_plain13b.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:pbatch -s 6 -c 1 -o gpurun_out/r2_pbatch_rx13 $CMD2 > gpurun_out/r2_ncu13b.log 2>&1
echo "ncu rx rc=$?"
CMD="python bench.py --steps 20 --warmup 5 --no-cpu --no-extras --e2e-sweeps 4"
$CMD > gpurun_out/r2_plain13.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:pbatch -s 2 -c 1 -o gpurun_out/r2_pbatch13 $CMD > gpurun_out/r2_ncu13.log 2>&1
echo "ncu single rc=$?"
