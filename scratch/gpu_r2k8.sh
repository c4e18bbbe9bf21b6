#!/bin/bash
# 8-GPU pass: exchange parity at world 8 x multimem 0 / 1, then the exchange modes at N = 8 (bench self-check included)
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r2k_gpus.txt
w=8
( time timeout 600 python -m pytest -v "tests/test_gpu_peer.py::test_peer_exchange_matches_nccl[$w-0]" "tests/test_gpu_peer.py::test_peer_exchange_matches_nccl[$w-1]" ) > gpurun_out/r2_peer_parity_w$w.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/r2_peer_parity_w$w.txt
grep -E "PASSED|FAILED|SKIPPED|passed|failed|rc=|real" gpurun_out/r2_peer_parity_w$w.txt | tail -6
run() { n=$1; tag=$2; shift; shift
  ( time env "$@" timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n --steps 40 --warmup 5 $ARGS ) > gpurun_out/r2k_n${n}_$tag.json 2> gpurun_out/r2k_n${n}_$tag.err
  python - gpurun_out/r2k_n${n}_$tag.json <<'P'
import json,sys
try:
    d=json.loads([l for l in open(sys.argv[1]).read().splitlines() if l.startswith("{")][0])
    c=d.get('exchange_check',{})
    print(sys.argv[1].split('r2k_')[1], d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], 'ident', d.get('replicas_identical'), 'close', c.get('frac_params_close'), 'dloss', c.get('loss_abs_diff'))
except Exception as e:
    print(sys.argv[1], 'ERR', e); print(open(sys.argv[1].replace('.json','.err')).read()[-1200:])
P
  grep real gpurun_out/r2k_n${n}_$tag.err
}
ARGS="" run 8 auto X=1
ARGS="--exchange hybrid" run 8 hybrid X=1
ARGS="--exchange peer" run 8 peer_nooverlap BRUSH_DP_OVERLAP=0
