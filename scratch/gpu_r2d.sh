#!/bin/bash
mkdir -p gpurun_out
( time python -m pytest tests -m gpu -q -x ) > gpurun_out/r2d_pytest.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/r2d_pytest.txt
python bench.py --steps 100 --warmup 5 --no-cpu-baseline > gpurun_out/r2d_bench.json 2> gpurun_out/r2d_bench.err
BRUSH_SPLIT_COLORS=1 python bench.py --steps 100 --warmup 5 --no-cpu-baseline > gpurun_out/r2d_bench_split.json 2> gpurun_out/r2d_bench_split.err
python bench.py --workload cfg5_6m_4k --capacity 167772160 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r2d_bench_cfg5.json 2> gpurun_out/r2d_bench_cfg5.err
python scratch/sortbench.py > gpurun_out/r2d_sort.txt 2>&1
for h in 148 296; do BRUSH_HIST_BLOCKS=$h python scratch/sortbench.py 2>&1 | head -4 > gpurun_out/r2d_sort_hist$h.txt; done
tail -4 gpurun_out/r2d_pytest.txt
for t in bench bench_split bench_cfg5; do python - gpurun_out/r2d_$t.json <<'P'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'])
    print({s['stage'][:14]: s['ms'] for s in d['stages']})
except Exception as e:
    print(sys.argv[1], 'ERR', e); print(open(sys.argv[1].replace('.json','.err')).read()[-800:])
P
done
cat gpurun_out/r2d_sort.txt | head -8; cat gpurun_out/r2d_sort_hist148.txt gpurun_out/r2d_sort_hist296.txt
