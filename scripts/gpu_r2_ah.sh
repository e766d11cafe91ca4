#!/bin/bash
# P2 convection matrix (non-symmetric contribution sums) + the other element types
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_p2.py tests/test_gpu_nedelec.py tests/test_gpu_2d.py tests/test_zz_no_alloc.py -m gpu -q -x > gpurun_out/r2ah_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2ah_pytest.log
tail -30 gpurun_out/r2ah_pytest.log
