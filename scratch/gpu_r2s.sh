#!/bin/bash
mkdir -p gpurun_out
for v in rs3 rs4; do
echo "== $v"
BRUSH_B200_LIB=$PWD/brush_b200/lib/libbrush_b200_$v.so bash scratch/gpu_q.sh $v
done
