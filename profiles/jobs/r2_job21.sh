#!/bin/bash
# Round 2, GPU job 21: whole GPU suite + smoke + default bench after the row-streaming kernel K13 and the pipelined end-to-end leg.
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/r2_pytest21.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2_pytest21.log
timeout 300 python __graft_entry__.py --smoke > gpurun_out/r2_smoke21.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/r2_smoke21.log
timeout 600 python bench.py > gpurun_out/r2_bench21_default.json 2> gpurun_out/r2_bench21_default.err; echo "bench default rc=$?"; tail -2 gpurun_out/r2_bench21_default.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2_bench21_default.json').read())
print('value',d['value'],'frac',d['roofline']['frac'])
print('e2e',d['e2e']['value'],d['e2e']['at_1_sweep_per_step'],d['e2e']['serial'])
print('one',d['one_sweep_per_launch'])
print('rx',d['rx']['value'],d['rx']['exchange_us'],d['rx']['parity'])
PY
