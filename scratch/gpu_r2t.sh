#!/bin/bash
mkdir -p gpurun_out
( python -m pytest tests/test_gpu_parity.py -m gpu -q -x ) > gpurun_out/t_pytest.txt 2>&1; tail -2 gpurun_out/t_pytest.txt
bash scratch/gpu_q.sh t
