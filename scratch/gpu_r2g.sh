#!/bin/bash
# 2-GPU: how thin must the side-stream SH exchange be to overlap with the next forward?
mkdir -p gpurun_out
run() { tag=$1; shift
  env "$@" python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 60 --warmup 5 --exchange peer > gpurun_out/r2g_$tag.json 2> gpurun_out/r2g_$tag.err
  python - gpurun_out/r2g_$tag.json <<'P'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1].split('r2g_')[1], d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], {s['stage'][:10]: s['ms'] for s in d['stages']})
except Exception as e:
    print(sys.argv[1], 'ERR', e)
P
}
run c1 BRUSH_DP_SIDE_CTAS=1
run c2 BRUSH_DP_SIDE_CTAS=2
run nooverlap BRUSH_DP_OVERLAP=0
