#!/bin/bash
# Round 2, GPU job 4: K12 with the incremental inner loop, the one-colour reduction and the leaner barrier.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/r2_pytest4.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/r2_pytest4.log
timeout 600 python tests/pbatch_probe.py --out gpurun_out/r2_pbatch_probe4.json > gpurun_out/r2_pbatch_probe4.log 2>&1; echo "probe rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2_pbatch_probe4.json'))
for L,row in d['single'].items(): print(L, {k:round(v.get('ms_per_sweep',-1)*1e3,2) for k,v in row.items()})
for R,row in d['rx'].items(): print(R, {k:round(v.get('ms_per_sweep',-1)*1e3,2) for k,v in row.items()})
print(d['latgas'])
PY
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu > gpurun_out/r2_bench4.json 2> gpurun_out/r2_bench4.err; echo "bench rc=$?"
cat gpurun_out/r2_bench4.json
