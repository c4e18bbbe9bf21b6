#!/bin/bash
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
run() { echo "== $*"; env "$@" timeout 200 $TR scratch/xchg_trace.py 2>&1 | grep "^step" | grep -v "step 8\|step 9\|step 12" ; }
( run X=1
run BRUSH_DP_OVERLAP=0
run BRUSH_DP_SIDE_CTAS=1
run BRUSH_DP_SIDE_BLOCKS=74
run BRUSH_DP_SIDE_BLOCKS=37
run BRUSH_DP_SIDE_BLOCKS=74 BRUSH_DP_U=8
run BRUSH_PDL=0 ) | tee gpurun_out/r2m_trace_n$N.txt
