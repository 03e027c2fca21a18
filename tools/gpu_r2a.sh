#!/bin/bash
# round 2, first visit: wide-kernel parity + resident-rate probes
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/r2a_smi.txt 2>&1
timeout 1500 python -m pytest tests/test_gpu_wide.py -m gpu -x -q 2>&1 | tail -30 > gpurun_out/r2a_pytest_wide.txt; tail -5 gpurun_out/r2a_pytest_wide.txt
timeout 600 python tools/wide_quick.py contract 592 > gpurun_out/r2a_quick_contract.txt 2>&1; tail -2 gpurun_out/r2a_quick_contract.txt
timeout 600 python tools/wide_quick.py long_runs_trim14 1184 > gpurun_out/r2a_quick_lr14.txt 2>&1; tail -2 gpurun_out/r2a_quick_lr14.txt
timeout 600 python tools/wide_quick.py long_runs_trim14 1184 SA_WIDE_ONLY=1 > gpurun_out/r2a_quick_lr14_wide.txt 2>&1; tail -2 gpurun_out/r2a_quick_lr14_wide.txt
timeout 600 python tools/wide_quick.py dense 2368 > gpurun_out/r2a_quick_dense.txt 2>&1; tail -2 gpurun_out/r2a_quick_dense.txt
timeout 600 python tools/wide_quick.py dense 2368 SA_WIDE_ONLY=1 > gpurun_out/r2a_quick_dense_wide.txt 2>&1; tail -2 gpurun_out/r2a_quick_dense_wide.txt
timeout 1500 python -m pytest tests -m gpu -x -q --deselect tests/test_gpu_wide.py 2>&1 | tail -15 > gpurun_out/r2a_pytest_rest.txt; tail -5 gpurun_out/r2a_pytest_rest.txt
