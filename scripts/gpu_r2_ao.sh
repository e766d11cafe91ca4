#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_edge_cases.py -m gpu -q > gpurun_out/r2ao_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2ao_pytest.log
grep -E "^E  |passed|failed|Error|^tests" gpurun_out/r2ao_pytest.log | head -40
