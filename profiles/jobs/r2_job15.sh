#!/bin/bash
# Round 2, GPU job 15: asynchronous upload/download + SweepPipeline tests; default bench with the pipelined end-to-end leg.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_pipeline.py -q -x > gpurun_out/r2_pytest15.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/r2_pytest15.log
timeout 600 python bench.py --steps 200 --warmup 20 --no-cpu > gpurun_out/r2_bench15.json 2> gpurun_out/r2_bench15.err; echo "bench rc=$?"; tail -3 gpurun_out/r2_bench15.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2_bench15.json').read())
print(d['value'], d['e2e'])
PY
for dep in 2 4; do timeout 600 python bench.py --steps 50 --warmup 5 --no-cpu --no-extras --e2e-depth $dep > gpurun_out/r2_bench15_d$dep.json 2>/dev/null; python -c "
import json; d=json.loads(open('gpurun_out/r2_bench15_d$dep.json').read()); print($dep, d['e2e']['value'], d['e2e']['at_1_sweep_per_step'], d['e2e']['serial'])"; done
