#!/bin/bash
N=${1:-2}
mkdir -p gpurun_out
( python -m pytest tests/test_gpu_train.py -m gpu -q -x -k "hybrid or views" ) > gpurun_out/u_pytest1.txt 2>&1; tail -2 gpurun_out/u_pytest1.txt
( timeout 300 python -m pytest -q "tests/test_gpu_peer.py::test_peer_exchange_matches_nccl[$N-1]" ) > gpurun_out/u_pytest2.txt 2>&1; tail -2 gpurun_out/u_pytest2.txt
for mode in hybrid auto; do
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 40 --warmup 5 --exchange $mode > gpurun_out/u_n${N}_$mode.json 2> gpurun_out/u_n${N}_$mode.err
python - gpurun_out/u_n${N}_$mode.json <<'P'
import json,sys
try:
    d=json.loads([l for l in open(sys.argv[1]).read().splitlines() if l.startswith("{")][0])
    c=d.get('exchange_check',{}); x=d.get('exchange_timing',{})
    print(sys.argv[1], d['value'], d['ms_per_step'], 'ident', d.get('replicas_identical'), 'close', c.get('frac_params_close'), 'xchg', x.get('exchange_ms'), x.get('to_exchange_ms_mean'))
except Exception as e:
    print('ERR', e); print(open(sys.argv[1].replace('.json','.err')).read()[-1500:])
P
done
