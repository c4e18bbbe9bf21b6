#!/bin/bash
# round-2 second GPU pass: parity suite (normal + poisoned buffers), sort A/B + phase clocks, sanitizer passes
mkdir -p gpurun_out
rm -f gpurun_out/parity_counts.json
( time python -m pytest tests -m gpu -q ) > gpurun_out/r2b_pytest.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/r2b_pytest.txt
cp gpurun_out/parity_counts.json gpurun_out/r2b_parity_counts.json 2>/dev/null
( BRUSH_POISON=1 BRUSH_PARITY_COUNTS=/tmp/pc.json python -m pytest tests -m gpu -q ) > gpurun_out/r2b_pytest_poison.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/r2b_pytest_poison.txt
python scratch/sortbench.py > gpurun_out/r2b_sort_items16.txt 2>&1
BRUSH_SORT_ITEMS=8 python scratch/sortbench.py > gpurun_out/r2b_sort_items8.txt 2>&1
BRUSH_B200_LIB=$PWD/brush_b200/lib/libbrush_b200_sortprof.so python scratch/sortbench.py > gpurun_out/r2b_sort_prof16.txt 2>&1
BRUSH_SORT_ITEMS=8 BRUSH_B200_LIB=$PWD/brush_b200/lib/libbrush_b200_sortprof.so python scratch/sortbench.py > gpurun_out/r2b_sort_prof8.txt 2>&1
python bench.py --steps 100 --warmup 5 --no-cpu-baseline > gpurun_out/r2b_bench.json 2> gpurun_out/r2b_bench.err
BRUSH_SORT_ITEMS=8 python bench.py --steps 100 --warmup 5 --no-cpu-baseline > gpurun_out/r2b_bench_items8.json 2> gpurun_out/r2b_bench_items8.err
SEL="test_forward_backward_vs_oracle and 5_000 or test_many_splats or test_trainer_step_reduces or test_views_step_two_views_against_oracle or test_sorting_big_vs_oracle and 100_003"
for tool in memcheck racecheck initcheck synccheck; do
  ( time timeout 600 compute-sanitizer --tool $tool --error-exitcode 9 python -m pytest tests/test_gpu_parity.py tests/test_gpu_train.py -q -x -k "test_many_splats or test_trainer_step_reduces or test_views_step_two_views_against_oracle or test_empty_and_single" ) > gpurun_out/r2b_sanitizer_$tool.txt 2>&1
  echo "rc=$?" >> gpurun_out/r2b_sanitizer_$tool.txt
done
tail -3 gpurun_out/r2b_pytest.txt gpurun_out/r2b_pytest_poison.txt
for t in memcheck racecheck initcheck synccheck; do tail -4 gpurun_out/r2b_sanitizer_$t.txt; done
