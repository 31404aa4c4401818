#!/bin/bash
# Round 2, GPU job 12 (8 GPUs): multi-GPU parity scripts, bench.py at N = 8 (config 3 strong + weak + in-job N = 1),
# the NCCL exchange path for comparison, config 5 (64 temperatures of the hard-core lattice gas over 8 GPUs; disc batch
# split over 8 GPUs), slab strong scaling.
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR tests/mgpu_rx_check.py > gpurun_out/r2_mgpu_rx12.log 2>&1; echo "mgpu rx rc=$?"; tail -8 gpurun_out/r2_mgpu_rx12.log
timeout 600 $TR bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/r2_bench12_n8.json 2> gpurun_out/r2_bench12_n8.err; echo "bench n8 rc=$?"; tail -2 gpurun_out/r2_bench12_n8.err
timeout 600 $TR bench.py --gpus 8 --steps 20 --warmup 5 --rx-comm nccl --no-extras > gpurun_out/r2_bench12_n8_nccl.json 2> gpurun_out/r2_bench12_n8_nccl.err; echo "bench n8 nccl rc=$?"
timeout 600 $TR bench.py --gpus 8 --steps 20 --warmup 5 --rx-every 1 --no-extras > gpurun_out/r2_bench12_n8_every1.json 2> gpurun_out/r2_bench12_n8_every1.err; echo "bench n8 every1 rc=$?"
timeout 600 $TR bench.py --gpus 8 --workload rx --rx-model hardcore --rx-replicas-total 64 --rx-kT-lo 0.5 --rx-kT-hi 2.0 --steps 20 --warmup 5 --no-extras > gpurun_out/r2_bench12_n8_hardcore.json 2> gpurun_out/r2_bench12_n8_hardcore.err; echo "bench n8 hardcore rc=$?"
timeout 600 $TR bench.py --gpus 8 --workload discs --steps 30 --no-cpu > gpurun_out/r2_bench12_n8_discs.json 2> gpurun_out/r2_bench12_n8_discs.err; echo "bench n8 discs rc=$?"
timeout 600 $TR bench.py --gpus 8 --workload slab --steps 200 --warmup 20 > gpurun_out/r2_bench12_n8_slab.json 2> gpurun_out/r2_bench12_n8_slab.err; echo "bench n8 slab rc=$?"
timeout 600 $TR tests/mgpu_slab_check.py > gpurun_out/r2_mgpu_slab12.log 2>&1; echo "mgpu slab rc=$?"; tail -3 gpurun_out/r2_mgpu_slab12.log
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_bench12_*.json')):
    try:
        d=json.loads([l for l in open(f).read().splitlines() if l.startswith('{')][-1])
        print(f.split('r2_bench12_')[1], '%.3e'%d['value'], 'ms/step', round(d['ms_per_step'],5), 'xus', d.get('exchange_us'), d.get('exchange_comm'), d.get('parity'), 'launches', d.get('gpu_launches'))
        if 'weak' in d: print('   weak', d['weak'])
        if 'scaling_check' in d: print('   check', d['scaling_check'])
    except Exception as e: print(f,'ERR',e)
PY
