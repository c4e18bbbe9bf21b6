#!/bin/bash
# 8-GPU pass: exchange parity tests at world 2 / 4 / 8 x multimem 0 / 1, then the exchange modes at N = 8 and N = 4
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r2i_gpus.txt
for w in 2 4 8; do
  ( time python -m pytest -v "tests/test_gpu_peer.py::test_peer_exchange_matches_nccl[$w-0]" "tests/test_gpu_peer.py::test_peer_exchange_matches_nccl[$w-1]" ) > gpurun_out/r2_peer_parity_w$w.txt 2>&1
  echo "pytest rc=$?" >> gpurun_out/r2_peer_parity_w$w.txt
  grep -E "PASSED|FAILED|SKIPPED|passed|failed|rc=" gpurun_out/r2_peer_parity_w$w.txt | tail -4
done
run() { n=$1; tag=$2; shift; shift
  env "$@" python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n --steps 40 --warmup 5 $ARGS > gpurun_out/r2i_n${n}_$tag.json 2> gpurun_out/r2i_n${n}_$tag.err
  python - gpurun_out/r2i_n${n}_$tag.json <<'P'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    c=d.get('exchange_check',{})
    print(sys.argv[1].split('r2i_')[1], d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], 'ident', d.get('replicas_identical'), 'close', c.get('frac_params_close'), 'dloss', c.get('loss_abs_diff'))
except Exception as e:
    print(sys.argv[1], 'ERR', e); print(open(sys.argv[1].replace('.json','.err')).read()[-1200:])
P
}
ARGS="--exchange hybrid" run 8 hybrid X=1
ARGS="--exchange peer" run 8 peer_overlap X=1
ARGS="--exchange peer" run 8 peer_nooverlap BRUSH_DP_OVERLAP=0
ARGS="--exchange hybrid" run 4 hybrid X=1
ARGS="--exchange views" run 4 views X=1
