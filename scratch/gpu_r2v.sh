#!/bin/bash
mkdir -p gpurun_out
for N in 8 4; do
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 40 --warmup 5 --exchange hybrid > gpurun_out/v_n${N}_hybrid.json 2> gpurun_out/v_n${N}_hybrid.err
python - gpurun_out/v_n${N}_hybrid.json <<'P'
import json,sys
try:
    d=json.loads([l for l in open(sys.argv[1]).read().splitlines() if l.startswith("{")][0])
    c=d.get('exchange_check',{}); x=d.get('exchange_timing',{})
    print(sys.argv[1], d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], 'ident', d.get('replicas_identical'), 'close', c.get('frac_params_close'), 'xchg', x.get('exchange_ms'), x.get('to_exchange_ms_mean'))
except Exception as e:
    print('ERR', e); print(open(sys.argv[1].replace('.json','.err')).read()[-1500:])
P
done
