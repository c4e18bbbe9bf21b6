#!/bin/bash
# 2-GPU: hybrid exchange parity + speed against the other modes
mkdir -p gpurun_out
( time python -m pytest tests/test_gpu_peer.py tests/test_gpu_train.py -q -x -k "peer or hybrid or views" ) > gpurun_out/r2h_pytest.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/r2h_pytest.txt
tail -5 gpurun_out/r2h_pytest.txt
run() { tag=$1; shift
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 60 --warmup 5 "$@" > gpurun_out/r2h_$tag.json 2> gpurun_out/r2h_$tag.err
  python - gpurun_out/r2h_$tag.json <<'P'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    c=d.get('exchange_check',{})
    print(sys.argv[1].split('r2h_')[1], d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], 'ident', d.get('replicas_identical'), 'close', c.get('frac_params_close'), 'dloss', c.get('loss_abs_diff'))
except Exception as e:
    print(sys.argv[1], 'ERR', e); print(open(sys.argv[1].replace('.json','.err')).read()[-1500:])
P
}
run hybrid --exchange hybrid
run views --exchange views
