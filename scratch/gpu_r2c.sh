#!/bin/bash
# 2-GPU pass: peer-exchange parity tests (world 2, multimem 0/1), overlap A/B, exchange modes
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r2c_gpus.txt
( time python -m pytest tests/test_gpu_peer.py tests/test_gpu_parity.py -q -k "peer or dp_step or deferred" ) > gpurun_out/r2c_pytest.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/r2c_pytest.txt
run() { # tag, extra env..., args
  tag=$1; shift
  env "$@" python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 60 --warmup 5 $ARGS > gpurun_out/r2c_bench_$tag.json 2> gpurun_out/r2c_bench_$tag.err
  tail -c 300 gpurun_out/r2c_bench_$tag.err | tail -2
}
ARGS="--exchange peer" run peer_overlap BRUSH_X=1
ARGS="--exchange peer" run peer_nooverlap BRUSH_DP_OVERLAP=0
ARGS="--exchange peer" run peer_overlap_mm BRUSH_DP_MULTIMEM=1
ARGS="--exchange peer" run peer_overlap_c4 BRUSH_DP_SIDE_CTAS=4
ARGS="--exchange views" run views BRUSH_X=1
tail -3 gpurun_out/r2c_pytest.txt
for t in peer_overlap peer_nooverlap peer_overlap_mm peer_overlap_c4 views; do python - gpurun_out/r2c_bench_$t.json <<'P'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    c=d.get('exchange_check',{})
    print(sys.argv[1].split('bench_')[1], d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], 'ident', d.get('replicas_identical'), 'maxrel', c.get('max_rel_param_diff'), 'close', c.get('frac_params_close'), 'dloss', c.get('loss_abs_diff'))
except Exception as e:
    print(sys.argv[1], 'ERR', e)
P
done
