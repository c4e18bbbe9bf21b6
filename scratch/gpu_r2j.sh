#!/bin/bash
# no-alloc scan test alone in a fresh process, then the whole GPU suite (it runs last there), then the default bench line
mkdir -p gpurun_out
( time python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x ) > gpurun_out/r2j_noalloc_alone.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/r2j_noalloc_alone.txt
tail -4 gpurun_out/r2j_noalloc_alone.txt
( time python -m pytest tests -m gpu -q -x --durations=8 ) > gpurun_out/r2j_pytest.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/r2j_pytest.txt
tail -16 gpurun_out/r2j_pytest.txt
( time python bench.py ) > gpurun_out/r2j_bench.json 2> gpurun_out/r2j_bench.err
python bench.py --workload cfg5_6m_4k --capacity 167772160 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r2j_bench_cfg5.json 2> gpurun_out/r2j_bench_cfg5.err
for t in bench bench_cfg5; do python - gpurun_out/r2j_$t.json <<'P'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[0])
    print(sys.argv[1], d['value'], d['ms_per_step'], 'e2e', d['e2e']['value'], 'launches', d.get('gpu_launches'))
    print({s['stage'][:14]: (s['ms'], s.get('frac_hbm')) for s in d['stages']})
    print(d['roofline'])
    print(d.get('cpu_baseline'), d.get('clocks'))
except Exception as e:
    print(sys.argv[1], 'ERR', e); print(open(sys.argv[1].replace('.json','.err')).read()[-800:])
P
done
