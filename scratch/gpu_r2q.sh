#!/bin/bash
mkdir -p gpurun_out
( python -m pytest tests/test_gpu_parity.py tests/test_gpu_train.py tests/test_abi_c.py -m gpu -q -x ) > gpurun_out/q_pytest.txt 2>&1; tail -3 gpurun_out/q_pytest.txt
python scratch/fwd_hash.py > gpurun_out/q_hash.txt 2>&1; diff gpurun_out/q_hash.txt gpurun_out/hash0.txt > /dev/null 2>&1; cat gpurun_out/q_hash.txt | head -3
NOCFG5= bash scratch/gpu_q.sh q
