#!/bin/bash
# Round 2, GPU job 39 (8 GPUs): final build at N = 8 -- multi-GPU parity (replica exchange, slabs), config 3 as written (+ weak,
# + in-job N = 1), exchange every sweep, config 5 as a lattice model, the slab workload.
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511"
timeout 300 $TR tests/mgpu_rx_check.py > gpurun_out/r2_mgpu_rx39.log 2>&1; echo "mgpu rx rc=$?"; grep MGPU gpurun_out/r2_mgpu_rx39.log
timeout 300 $TR tests/mgpu_slab_check.py > gpurun_out/r2_mgpu_slab39.log 2>&1; echo "mgpu slab rc=$?"; grep MGPU gpurun_out/r2_mgpu_slab39.log
timeout 600 $TR bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/r2_bench39_n8.json 2> gpurun_out/r2_bench39_n8.err; echo "bench n8 rc=$?"
timeout 600 $TR bench.py --gpus 8 --steps 20 --warmup 5 --rx-every 1 --no-extras > gpurun_out/r2_bench39_n8_every1.json 2> /dev/null; echo "bench n8 every1 rc=$?"
timeout 600 $TR bench.py --gpus 8 --workload rx --rx-model hardcore --rx-replicas-total 64 --rx-kT-lo 0.5 --rx-kT-hi 2.0 --steps 20 --warmup 5 --no-extras > gpurun_out/r2_bench39_n8_hardcore.json 2> /dev/null; echo "bench n8 hardcore rc=$?"
timeout 600 $TR bench.py --gpus 8 --workload slab --steps 200 --warmup 20 > gpurun_out/r2_bench39_n8_slab.json 2> /dev/null; echo "bench n8 slab rc=$?"
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_bench39_*.json')):
    try:
        d=json.loads([l for l in open(f).read().splitlines() if l.startswith('{')][-1])
        print(f.split('r2_bench39_')[1], '%.3e'%d['value'], 'ms/step', round(d['ms_per_step'],5), 'xus', d.get('exchange_us'), d.get('parity'), 'e2e', d.get('e2e',{}).get('value'))
        if 'weak' in d: print('   weak %.4e'%d['weak']['value'], d['weak']['exchange_us'])
        if 'scaling_check' in d: print('   check', d['scaling_check']['efficiency'], d['scaling_check']['weak_efficiency'], '%.4e'%d['scaling_check']['n1_value'])
    except Exception as e: print(f, 'ERR', e)
PY
