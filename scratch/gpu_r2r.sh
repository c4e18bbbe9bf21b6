#!/bin/bash
# views kernel change: single-GPU-visible tests of the views / hybrid step on GPU 0, then world-2 parity + bench
mkdir -p gpurun_out
( python -m pytest tests/test_gpu_train.py -m gpu -q -x -k "views or hybrid or trainer" ) > gpurun_out/r_pytest1.txt 2>&1; tail -2 gpurun_out/r_pytest1.txt
( timeout 300 python -m pytest -q "tests/test_gpu_peer.py::test_peer_exchange_matches_nccl[2-0]" ) > gpurun_out/r_pytest2.txt 2>&1; tail -2 gpurun_out/r_pytest2.txt
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 40 --warmup 5 > gpurun_out/r_n2.json 2> gpurun_out/r_n2.err
python - gpurun_out/r_n2.json <<'P'
import json,sys
try:
    d=json.loads([l for l in open(sys.argv[1]).read().splitlines() if l.startswith("{")][0])
    c=d.get('exchange_check',{}); x=d.get('exchange_timing',{})
    print(d['value'], d['ms_per_step'], 'ident', d.get('replicas_identical'), 'close', c.get('frac_params_close'), 'xchg', x.get('exchange_ms'), x.get('to_exchange_ms_mean'))
except Exception as e:
    print('ERR', e); print(open(sys.argv[1].replace('.json','.err')).read()[-1500:])
P
