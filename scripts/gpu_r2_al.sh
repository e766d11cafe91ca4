#!/bin/bash
# demagnetising field on partitioned handles (fk_multi_demag_*), 2 GPUs
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_multi.py -m gpu -q -x -k "demag or single_process" > gpurun_out/r2al_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2al_pytest.log
tail -30 gpurun_out/r2al_pytest.log
