#!/bin/bash
# Round 2, GPU job 7: K12 with export rows first and incremental row tracking, per-block exchange decisions: tests, probe, ncu.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_zx_gpu_pbatch.py tests/test_zy_gpu_persist.py tests/test_gpu_resident.py -q > gpurun_out/r2_pytest7.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/r2_pytest7.log
timeout 600 python tests/pbatch_probe.py --out gpurun_out/r2_pbatch_probe7.json > gpurun_out/r2_pbatch_probe7.log 2>&1; echo "probe rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2_pbatch_probe7.json'))
for L,row in d['single'].items(): print(L, {k:round(v.get('ms_per_sweep',-1)*1e3,2) for k,v in row.items()})
for R,row in d['rx'].items(): print(R, {k:round(v.get('ms_per_sweep',-1)*1e3,2) for k,v in row.items()})
print(d['latgas'])
PY
CMD="python bench.py --steps 20 --warmup 5 --no-cpu --no-extras --e2e-sweeps 4"
$CMD > gpurun_out/r2_plain7.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:pbatch -s 2 -c 1 -o gpurun_out/r2_pbatch7 $CMD > gpurun_out/r2_ncu7.log 2>&1
echo "ncu single rc=$?"
