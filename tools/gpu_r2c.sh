#!/bin/bash
# wide-kernel iteration: parity (wide tests + logAdd bit test) then rate probes
mkdir -p gpurun_out
T=${1:-r2c}
timeout 1500 python -m pytest tests/test_gpu_wide.py tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -30 > gpurun_out/${T}_pytest.txt; tail -4 gpurun_out/${T}_pytest.txt
timeout 600 python tools/wide_quick.py contract 1184 > gpurun_out/${T}_quick_contract.txt 2>&1; tail -1 gpurun_out/${T}_quick_contract.txt | cut -c1-900
timeout 600 python tools/wide_quick.py dense 2368 SA_WIDE_ONLY=1 > gpurun_out/${T}_quick_dense_wide.txt 2>&1; tail -1 gpurun_out/${T}_quick_dense_wide.txt | cut -c1-600
timeout 600 python tools/wide_quick.py dense 2368 > gpurun_out/${T}_quick_dense.txt 2>&1; tail -1 gpurun_out/${T}_quick_dense.txt | cut -c1-600
timeout 600 python tools/wide_quick.py long_runs_trim14 1184 SA_WIDE_ONLY=1 > gpurun_out/${T}_quick_lr14_wide.txt 2>&1; tail -1 gpurun_out/${T}_quick_lr14_wide.txt | cut -c1-600
