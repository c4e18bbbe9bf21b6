#!/bin/bash
# usage: scratch/quickbench.sh TAG  -> runs gpu tests + bench, prints stage table
TAG=${1:-x}
python -m pytest tests -m gpu -x -q 2>&1 | tail -40 > gpurun_out/pytest_$TAG.log
python bench.py --steps 30 --warmup 5 --no-cpu-baseline > gpurun_out/bench_$TAG.log 2>&1
tail -15 gpurun_out/pytest_$TAG.log
python -c "
import json
d=json.loads(open('gpurun_out/bench_$TAG.log').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['e2e']['value'])
print({s['stage']: s['ms'] for s in d['stages']})
"
