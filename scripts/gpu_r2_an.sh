#!/bin/bash
# final state on a 2-GPU box: the whole GPU suite (multi-GPU tests included)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2an_pytest_gpu_2gpu.log 2>&1; echo "pytest -m gpu rc=$?" | tee -a gpurun_out/r2an_pytest_gpu_2gpu.log
tail -5 gpurun_out/r2an_pytest_gpu_2gpu.log
