#!/bin/bash
# usage: scratch/gpu_quick.sh TAG  -> gpu tests + bench (cfg4 and cfg5), prints stage tables
TAG=${1:-x}
mkdir -p gpurun_out
( time python -m pytest tests -m gpu -q -x ) > gpurun_out/${TAG}_pytest.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/${TAG}_pytest.txt
python bench.py --steps 100 --warmup 5 --no-cpu-baseline > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
python bench.py --workload cfg5_6m_4k --capacity 167772160 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_bench_cfg5.json 2> gpurun_out/${TAG}_bench_cfg5.err
tail -4 gpurun_out/${TAG}_pytest.txt
for t in bench bench_cfg5; do python - gpurun_out/${TAG}_$t.json <<'P'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'])
    print({s['stage'][:14]: s['ms'] for s in d['stages']})
    print({k:v for k,v in d['roofline'].items() if k in ('frac','frac_executed','frac_mufu','culled_pair_ratio','peak')})
except Exception as e:
    print(sys.argv[1], 'ERR', e); print(open(sys.argv[1].replace('.json','.err')).read()[-800:])
P
done
