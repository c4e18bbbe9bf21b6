#!/bin/bash
# final round-2 package: bench lines (ours, reference arm, fused-64 A/B at cfg4 and cfg5), ncu launch list, --set full of one step,
# executed-instruction counts of the rasterizers, --set full of the packed forward variant
TAG=r2p
mkdir -p gpurun_out
python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${TAG}_bench_reference.json 2> gpurun_out/${TAG}_bench_reference.err
python bench.py --steps 60 --sort-mode fused64 --no-cpu-baseline > gpurun_out/${TAG}_bench_fused64.json 2> gpurun_out/${TAG}_bench_fused64.err
python bench.py --workload cfg5_6m_4k --capacity 167772160 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_bench_cfg5.json 2> gpurun_out/${TAG}_bench_cfg5.err
python bench.py --workload cfg5_6m_4k --capacity 167772160 --steps 20 --warmup 3 --no-cpu-baseline --sort-mode fused64 > gpurun_out/${TAG}_bench_cfg5_fused64.json 2> gpurun_out/${TAG}_bench_cfg5_fused64.err
CMD="python bench.py --steps 2 --warmup 3 --profile-step"
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu1.log 2>&1
echo "ncu1 rc=$?"
ncu --set full --clock-control none --import-source on --profile-from-start off -f -o gpurun_out/${TAG}_step $CMD > gpurun_out/${TAG}_ncu2.log 2>&1
echo "ncu2 rc=$?"
ncu --metrics smsp__sass_thread_inst_executed_op_ffma_pred_on.sum,smsp__sass_thread_inst_executed_op_fmul_pred_on.sum,smsp__sass_thread_inst_executed_op_fadd_pred_on.sum,smsp__thread_inst_executed_pipe_xu_pred_on.sum,smsp__inst_executed.sum,smsp__thread_inst_executed.sum,gpu__time_duration.sum --clock-control none --profile-from-start off -k regex:rasterize --csv --log-file gpurun_out/${TAG}_executed.csv $CMD > gpurun_out/${TAG}_ncu3.log 2>&1
echo "ncu3 rc=$?"
BRUSH_RASTER_PACKED=1 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:rasterize_kernel -f -o gpurun_out/${TAG}_packed_fwd $CMD > gpurun_out/${TAG}_ncu4.log 2>&1
echo "ncu4 rc=$?"
for t in bench bench_fused64 bench_cfg5 bench_cfg5_fused64; do python - gpurun_out/${TAG}_$t.json <<'P'
import json,sys
try:
    d=json.loads([l for l in open(sys.argv[1]).read().splitlines() if l.startswith("{")][0])
    print(sys.argv[1], d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], {s['stage'][:10]: s['ms'] for s in d['stages']})
except Exception as e:
    print(sys.argv[1], 'ERR', e); print(open(sys.argv[1].replace('.json','.err')).read()[-800:])
P
done
tail -c 500 gpurun_out/${TAG}_bench_reference.json
ls -la gpurun_out | grep $TAG
