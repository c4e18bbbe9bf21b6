#!/bin/bash
# peaks + ncu: launch list of one steady-state step, --set full of it, executed FP32 instruction counts of the rasterizers
mkdir -p gpurun_out
python -c "
import torch
from brush_b200 import get_context
print(get_context('cuda:0').measure_peaks())
" > gpurun_out/r2e_peaks.txt 2>&1
cat gpurun_out/r2e_peaks.txt
CMD="python bench.py --steps 2 --warmup 3 --profile-step"
$CMD > gpurun_out/r2e_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/r2e_launches.csv $CMD > gpurun_out/r2e_ncu1.log 2>&1
echo "ncu1 rc=$?"
ncu --set full --clock-control none --import-source on --profile-from-start off -f -o gpurun_out/r2e_step $CMD > gpurun_out/r2e_ncu2.log 2>&1
echo "ncu2 rc=$?"
ncu --metrics smsp__sass_thread_inst_executed_op_ffma_pred_on.sum,smsp__sass_thread_inst_executed_op_fmul_pred_on.sum,smsp__sass_thread_inst_executed_op_fadd_pred_on.sum,smsp__thread_inst_executed_pipe_xu_pred_on.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum,gpu__time_duration.sum --clock-control none --profile-from-start off -k regex:rasterize --csv --log-file gpurun_out/r2e_executed.csv $CMD > gpurun_out/r2e_ncu3.log 2>&1
echo "ncu3 rc=$?"
ls -la gpurun_out | grep r2e
