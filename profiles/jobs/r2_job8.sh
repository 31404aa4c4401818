#!/bin/bash
# Round 2, GPU job 8: Philox round keys precomputed on the host (constant memory) in every kernel; whole GPU suite,
# probe, ncu of K12 and of the bit-packed row kernel.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/r2_pytest8.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/r2_pytest8.log
timeout 600 python tests/pbatch_probe.py --out gpurun_out/r2_pbatch_probe8.json > gpurun_out/r2_pbatch_probe8.log 2>&1; echo "probe rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2_pbatch_probe8.json'))
for L,row in d['single'].items(): print(L, {k:round(v.get('ms_per_sweep',-1)*1e3,2) for k,v in row.items()})
for R,row in d['rx'].items(): print(R, {k:round(v.get('ms_per_sweep',-1)*1e3,2) for k,v in row.items()})
print(d['latgas'])
PY
CMD="python bench.py --steps 20 --warmup 5 --no-cpu --no-extras --e2e-sweeps 4"
$CMD > gpurun_out/r2_plain8.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:pbatch -s 2 -c 1 -o gpurun_out/r2_pbatch8 $CMD > gpurun_out/r2_ncu8.log 2>&1
echo "ncu single rc=$?"
