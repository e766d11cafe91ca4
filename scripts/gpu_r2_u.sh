#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_p2.py tests/test_gpu_2d.py -m gpu -q -s -x > gpurun_out/r2u_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2u_pytest.log
grep -v "^$" gpurun_out/r2u_pytest.log | tail -25
