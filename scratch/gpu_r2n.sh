#!/bin/bash
mkdir -p gpurun_out
BRUSH_B200_LIB=$PWD/brush_b200/lib/libbrush_b200_slow.so python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r2n_slow.json 2> gpurun_out/r2n_slow.err
python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r2n_norm.json 2> gpurun_out/r2n_norm.err
python - <<'P'
import json
for t in ("slow","norm"):
    try:
        d=json.loads([l for l in open(f"gpurun_out/r2n_{t}.json").read().splitlines() if l.startswith("{")][0])
        s=[x for x in d['stages'] if x['stage']=='Rasterize'][0]
        print(t, 'fwd groups (warp-level):', s['pairs_evaluated']/128, 'pairs_ref', s['pairs_reference'], 'ms', s['ms'])
    except Exception as e:
        print(t, 'ERR', e, open(f"gpurun_out/r2n_{t}.err").read()[-600:])
P
