#!/bin/bash
mkdir -p gpurun_out
BRUSH_RASTER_PACKED=0 python scratch/fwd_hash.py > gpurun_out/hash0.txt 2>&1
BRUSH_RASTER_PACKED=1 python scratch/fwd_hash.py > gpurun_out/hash1.txt 2>&1
diff gpurun_out/hash0.txt gpurun_out/hash1.txt && echo "IDENTICAL BITS" || { echo DIFFER; tail -5 gpurun_out/hash1.txt; }
( python -m pytest tests/test_gpu_parity.py -m gpu -q -x ) > gpurun_out/p1_pytest.txt 2>&1; tail -3 gpurun_out/p1_pytest.txt
for pk in 0 1; do
BRUSH_RASTER_PACKED=$pk python bench.py --steps 60 --warmup 5 --no-cpu-baseline > gpurun_out/p1_bench_$pk.json 2> gpurun_out/p1_bench_$pk.err
python - gpurun_out/p1_bench_$pk.json <<'P'
import json,sys
try:
    d=json.loads([l for l in open(sys.argv[1]).read().splitlines() if l.startswith("{")][0])
    print(sys.argv[1], d['value'], d['ms_per_step'], {s['stage'][:10]: s['ms'] for s in d['stages']})
except Exception as e:
    print(sys.argv[1], 'ERR', e); print(open(sys.argv[1].replace('.json','.err')).read()[-800:])
P
done
BRUSH_RASTER_PACKED=1 python bench.py --workload cfg5_6m_4k --capacity 167772160 --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads([l for l in sys.stdin.read().splitlines() if l.startswith('{')][0]); print('cfg5', d['value'], {s['stage'][:10]: s['ms'] for s in d['stages']})"
