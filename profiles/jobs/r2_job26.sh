#!/bin/bash
# Round 2, GPU job 26 (8 GPUs): parity and bench at N = 8 after the packet exchange in K12: config 3 as written (+ weak,
# + in-job N = 1), exchange every sweep, config 5 as a lattice model (64 temperatures, 8 per GPU).
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR tests/mgpu_rx_check.py > gpurun_out/r2_mgpu_rx26.log 2>&1; echo "mgpu rx rc=$?"; tail -3 gpurun_out/r2_mgpu_rx26.log
timeout 600 $TR bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/r2_bench26_n8.json 2> gpurun_out/r2_bench26_n8.err; echo "bench n8 rc=$?"; tail -2 gpurun_out/r2_bench26_n8.err
timeout 600 $TR bench.py --gpus 8 --steps 20 --warmup 5 --rx-every 1 --no-extras > gpurun_out/r2_bench26_n8_every1.json 2> gpurun_out/r2_bench26_n8_every1.err; echo "bench n8 every1 rc=$?"
timeout 600 $TR bench.py --gpus 8 --workload rx --rx-model hardcore --rx-replicas-total 64 --rx-kT-lo 0.5 --rx-kT-hi 2.0 --steps 20 --warmup 5 --no-extras > gpurun_out/r2_bench26_n8_hardcore.json 2> gpurun_out/r2_bench26_n8_hardcore.err; echo "bench n8 hardcore rc=$?"
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_bench26_*.json')):
    try:
        d=json.loads([l for l in open(f).read().splitlines() if l.startswith('{')][-1])
        print(f.split('r2_bench26_')[1], '%.3e'%d['value'], 'ms/step', round(d['ms_per_step'],5), 'xus', d.get('exchange_us'), d.get('exchange_comm'), d.get('parity'), 'e2e %.3e'%d['e2e']['value'])
        if 'weak' in d: print('   weak', d['weak'])
        if 'scaling_check' in d: print('   check', d['scaling_check'])
    except Exception as e: print(f, 'ERR', e)
PY
