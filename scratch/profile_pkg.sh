#!/bin/bash
# usage: scratch/profile_pkg.sh TAG  -> bench line, reference line, ncu launch list, ncu --set full of one step
TAG=${1:-r1x}
python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${TAG}_bench_reference.json 2> gpurun_out/${TAG}_bench_reference.err
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/${TAG}_launches.csv \
    python bench.py --steps 2 --warmup 3 --profile-step > gpurun_out/${TAG}_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on --profile-from-start off -f -o gpurun_out/${TAG}_step \
    python bench.py --steps 2 --warmup 3 --profile-step > gpurun_out/${TAG}_ncu2.log 2>&1
tail -c 600 gpurun_out/${TAG}_bench.json; echo; tail -c 400 gpurun_out/${TAG}_bench_reference.json; ls -la gpurun_out/ | grep $TAG
