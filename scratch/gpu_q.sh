#!/bin/bash
# quick A/B: cfg4 + cfg5 bench lines (stage tables), optional pytest of the parity suite first when PYTEST=1
TAG=${1:-q}
mkdir -p gpurun_out
if [ -n "$PYTEST" ]; then ( python -m pytest tests -m gpu -q -x $PYTEST_ARGS ) > gpurun_out/${TAG}_pytest.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/${TAG}_pytest.txt; tail -3 gpurun_out/${TAG}_pytest.txt; fi
python bench.py --steps 60 --warmup 5 --no-cpu-baseline > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
[ -n "$NOCFG5" ] || python bench.py --workload cfg5_6m_4k --capacity 167772160 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_bench_cfg5.json 2> gpurun_out/${TAG}_bench_cfg5.err
for t in bench bench_cfg5; do [ -f gpurun_out/${TAG}_$t.json ] && python - gpurun_out/${TAG}_$t.json <<'P'
import json,sys
try:
    d=json.loads([l for l in open(sys.argv[1]).read().splitlines() if l.startswith("{")][0])
    print(sys.argv[1], d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'])
    print({s['stage'][:10]: s['ms'] for s in d['stages']})
except Exception as e:
    print(sys.argv[1], 'ERR', e); print(open(sys.argv[1].replace('.json','.err')).read()[-800:])
P
done
