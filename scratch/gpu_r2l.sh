#!/bin/bash
# N-GPU: exchange kernel alone, CTAs per SM x float4s in flight per thread; then the bench with exchange timing
N=${1:-8}
mkdir -p gpurun_out
for u in 2 4 8; do
  BRUSH_DP_U=$u timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 scratch/dpsweep.py 2>&1 | grep "world=" | tee -a gpurun_out/r2l_dpsweep_n$N.txt
done
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 40 --warmup 5 --exchange peer > gpurun_out/r2l_n${N}_peer.json 2> gpurun_out/r2l_n${N}_peer.err
python - gpurun_out/r2l_n${N}_peer.json <<'P'
import json,sys
d=json.loads([l for l in open(sys.argv[1]).read().splitlines() if l.startswith("{")][0])
print(d['value'], d['ms_per_step'], json.dumps(d.get('exchange_timing'))[:700])
P
