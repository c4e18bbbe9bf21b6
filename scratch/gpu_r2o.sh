#!/bin/bash
mkdir -p gpurun_out
for it in 16 8; do
BRUSH_SORT_ITEMS=$it python bench.py --workload cfg5_6m_4k --capacity 167772160 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2o_cfg5_items$it.json 2> gpurun_out/r2o_cfg5_items$it.err
python - gpurun_out/r2o_cfg5_items$it.json <<'P'
import json,sys
d=json.loads([l for l in open(sys.argv[1]).read().splitlines() if l.startswith("{")][0])
print(sys.argv[1], d['value'], d['ms_per_step'], {s['stage'][:10]: s['ms'] for s in d['stages']})
P
done
