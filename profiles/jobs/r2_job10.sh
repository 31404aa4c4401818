#!/bin/bash
# Round 2, GPU job 10 (2 GPUs): replica exchange through peer memory inside the sweep kernel -- parity script, then
# bench.py at N = 2 with both exchange paths.  Single-GPU tests of the changed exchange code first.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zx_gpu_pbatch.py tests/test_gpu_rx.py tests/test_gpu_rx_api.py -q > gpurun_out/r2_pytest10.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_pytest10.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR tests/mgpu_rx_check.py > gpurun_out/r2_mgpu_rx10.log 2>&1; echo "mgpu rx rc=$?"; tail -8 gpurun_out/r2_mgpu_rx10.log
timeout 600 $TR bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r2_bench10_peer.json 2> gpurun_out/r2_bench10_peer.err; echo "bench peer rc=$?"
tail -3 gpurun_out/r2_bench10_peer.err
timeout 600 $TR bench.py --gpus 2 --steps 20 --warmup 5 --rx-comm nccl --no-extras > gpurun_out/r2_bench10_nccl.json 2> gpurun_out/r2_bench10_nccl.err; echo "bench nccl rc=$?"
python - <<'PY'
import json
for f in ('peer','nccl'):
    try:
        d=json.loads(open('gpurun_out/r2_bench10_%s.json'%f).read())
        print(f, d['value'], d['ms_per_step'], d['exchange_us'], d.get('exchange_comm'), d['parity'], d.get('weak',{}).get('value'), d.get('scaling_check'))
    except Exception as e: print(f, 'ERR', e)
PY
timeout 300 python tests/pbatch_probe.py --out gpurun_out/r2_pbatch_probe10.json > gpurun_out/r2_pbatch_probe10.log 2>&1
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2_pbatch_probe10.json'))
for R,row in d['rx'].items(): print(R, {k:round(v.get('ms_per_sweep',-1)*1e3,2) for k,v in row.items() if k.startswith('pbatch')})
PY
