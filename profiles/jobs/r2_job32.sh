#!/bin/bash
# Round 2, GPU job 32 (8 GPUs): one 16384^2 lattice in row slabs over 4 and 8 GPUs with the slabs on chip (K12 + packets over
# NVLink): parity check and strong-scaling bench; the default replica-exchange bench at N = 4 (not run before this round).
mkdir -p gpurun_out
for N in 8 4; do
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
timeout 300 $TR tests/mgpu_slab_check.py > gpurun_out/r2_mgpu_slab32_$N.log 2>&1; echo "mgpu slab $N rc=$?"; grep "slab\|MGPU" gpurun_out/r2_mgpu_slab32_$N.log | tail -6
timeout 300 $TR bench.py --gpus $N --workload slab --steps 200 --warmup 20 > gpurun_out/r2_bench32_n${N}_slab.json 2> gpurun_out/r2_bench32_n${N}_slab.err; echo "bench slab $N rc=$?"; cut -c1-260 gpurun_out/r2_bench32_n${N}_slab.json
done
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR bench.py --gpus 4 --steps 20 --warmup 5 > gpurun_out/r2_bench32_n4.json 2> gpurun_out/r2_bench32_n4.err; echo "bench n4 rc=$?"
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r2_bench32_n4.json') if l.startswith('{')][-1])
print('n4 %.3e'%d['value'], d['ms_per_step'], 'xus', d.get('exchange_us'), d['parity'], 'e2e %.3e'%d['e2e']['value'], d.get('scaling_check',{}).get('efficiency'), d.get('weak',{}).get('value'))
PY
