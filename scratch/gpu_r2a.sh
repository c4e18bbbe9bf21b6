#!/bin/bash
# round-2 first GPU pass: parity suite, bench with PDL on/off, config 5, reference arm numbers
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.max.mem --format=csv > gpurun_out/r2a_gpu.txt
( time python -m pytest tests -m gpu -x -q ) > gpurun_out/r2a_pytest.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/r2a_pytest.txt
python bench.py --steps 100 --warmup 5 > gpurun_out/r2a_bench.json 2> gpurun_out/r2a_bench.err
BRUSH_PDL=0 python bench.py --steps 100 --warmup 5 --no-cpu-baseline > gpurun_out/r2a_bench_nopdl.json 2> gpurun_out/r2a_bench_nopdl.err
python bench.py --workload cfg5_6m_4k --capacity 167772160 --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r2a_bench_cfg5.json 2> gpurun_out/r2a_bench_cfg5.err
python bench.py --impl reference --workload cfg1_16k_256 --capacity 200000 --steps 20 --warmup 2 > gpurun_out/r2a_ref_cfg1.json 2>&1
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2a_ref_cfg4.json 2>&1
nproc > gpurun_out/r2a_nproc.txt
tail -3 gpurun_out/r2a_pytest.txt
head -c 600 gpurun_out/r2a_bench.json
