#!/bin/bash
# Round 2, GPU job 25 (2 GPUs): multi-GPU parity and bench after the packet exchange in K12 and the asynchronous uploads.
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR tests/mgpu_rx_check.py > gpurun_out/r2_mgpu_rx25.log 2>&1; echo "mgpu rx rc=$?"; tail -6 gpurun_out/r2_mgpu_rx25.log
timeout 600 $TR tests/mgpu_slab_check.py > gpurun_out/r2_mgpu_slab25.log 2>&1; echo "mgpu slab rc=$?"; tail -3 gpurun_out/r2_mgpu_slab25.log
timeout 600 $TR bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r2_bench25_n2.json 2> gpurun_out/r2_bench25_n2.err; echo "bench n2 rc=$?"; tail -3 gpurun_out/r2_bench25_n2.err
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r2_bench25_n2.json') if l.startswith('{')][-1])
print('%.3e'%d['value'], d['ms_per_step'], 'xus', d.get('exchange_us'), d.get('exchange_comm'), d['parity'], 'e2e %.3e'%d['e2e']['value'], d['e2e'].get('at_1_sweep_per_step'))
print(' weak', d.get('weak')); print(' check', d.get('scaling_check'))
PY
