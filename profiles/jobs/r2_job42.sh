#!/bin/bash
# Round 2, GPU job 42 (N GPUs, $1): replica-exchange bench with the sums of an exchange travelling as packets, exchange every 10 / every 1.
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
for E in 10 1; do
  X=""; [ $E = 1 ] && X="--no-extras"
  timeout 600 $TR bench.py --gpus $N --steps 20 --warmup 5 --rx-every $E $X > gpurun_out/r2_bench42_n${N}_every$E.json 2> gpurun_out/r2_bench42_n${N}_every$E.err; echo "bench n$N every $E rc=$?"
done
python - <<PY
import json,glob
for f in sorted(glob.glob('gpurun_out/r2_bench42_n${N}_*.json')):
    try:
        d=json.loads([l for l in open(f).read().splitlines() if l.startswith('{')][-1])
        print(f.split('r2_bench42_')[1], '%.4e'%d['value'], 'ms/step', round(d['ms_per_step'],5), 'xus', d.get('exchange_us'), d.get('parity'), 'e2e', d.get('e2e',{}).get('value'))
        if 'weak' in d: print('   weak %.4e'%d['weak']['value'], d['weak']['exchange_us'])
        if 'scaling_check' in d: print('   check', d['scaling_check']['efficiency'], d['scaling_check']['weak_efficiency'], '%.4e'%d['scaling_check']['n1_value'])
    except Exception as e: print(f, 'ERR', e)
PY
