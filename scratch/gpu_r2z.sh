#!/bin/bash
# final check of the round: whole GPU suite, default bench line, launch list + --set full of one step with the final kernels
TAG=r2z
mkdir -p gpurun_out
( time python -m pytest tests -m gpu -q -x ) > gpurun_out/${TAG}_pytest.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/${TAG}_pytest.txt
tail -5 gpurun_out/${TAG}_pytest.txt
python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
CMD="python bench.py --steps 2 --warmup 3 --profile-step"
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu1.log 2>&1
echo "ncu1 rc=$?"
ncu --set full --clock-control none --import-source on --profile-from-start off -f -o gpurun_out/${TAG}_step $CMD > gpurun_out/${TAG}_ncu2.log 2>&1
echo "ncu2 rc=$?"
python - gpurun_out/${TAG}_bench.json <<'P'
import json,sys
try:
    d=json.loads([l for l in open(sys.argv[1]).read().splitlines() if l.startswith("{")][0])
    print(d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], {s['stage'][:10]: s['ms'] for s in d['stages']})
    print(d['roofline'].get('frac'), d['roofline'].get('frac_executed'), d['cpu_baseline']['value'], d['clocks'])
except Exception as e:
    print('ERR', e); print(open(sys.argv[1].replace('.json','.err')).read()[-800:])
P
