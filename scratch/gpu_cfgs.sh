#!/bin/bash
mkdir -p gpurun_out
for w in cfg1_16k_256 cfg2_100k_1024 cfg3_300k_800; do
python bench.py --workload $w --steps 100 --warmup 5 --no-cpu-baseline > gpurun_out/r2y_bench_$w.json 2> gpurun_out/r2y_bench_$w.err
python - gpurun_out/r2y_bench_$w.json <<'P'
import json,sys
try:
    d=json.loads([l for l in open(sys.argv[1]).read().splitlines() if l.startswith("{")][0])
    print(sys.argv[1].split('bench_')[1], d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], d['config']['num_visible'], d['config']['num_intersections'])
except Exception as e:
    print(sys.argv[1], 'ERR', e); print(open(sys.argv[1].replace('.json','.err')).read()[-600:])
P
done
