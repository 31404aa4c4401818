#!/bin/bash
# Round 2, GPU job 22: K12 with the exported rows travelling as {data, pass number} packets (no fence, no flag): parity
# tests of the batch-resident kernel and of replica exchange, then the probe over sizes / batches.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_zx_gpu_pbatch.py tests/test_gpu_rx.py tests/test_gpu_rx_api.py tests/test_gpu_errors.py tests/test_gpu_stats.py -q -x > gpurun_out/r2_pytest22.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_pytest22.log
timeout 600 python tests/pbatch_probe.py --out gpurun_out/r2_pbatch_probe22.json > gpurun_out/r2_pbatch_probe22.log 2>&1; echo "probe rc=$?"
python - <<'PY'
import json
new=json.load(open('gpurun_out/r2_pbatch_probe22.json')); old=json.load(open('profiles/r2_pbatch_probe.json'))
for name,d in (('old',old),('new',new)):
    print(name)
    for L,row in d['single'].items(): print('  single',L, {k:round(v.get('ms_per_sweep',-1)*1e3,2) for k,v in row.items() if isinstance(v,dict) and ('pbatch' in k)})
    for R,row in d['rx'].items(): print('  rx',R, {k:round(v.get('ms_per_sweep',-1)*1e3,2) for k,v in row.items() if isinstance(v,dict) and k.startswith('pbatch')})
    print('  latgas', d.get('latgas'))
PY
