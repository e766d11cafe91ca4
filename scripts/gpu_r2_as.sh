#!/bin/bash
mkdir -p gpurun_out
timeout 100 python -m pytest tests/test_gpu_nedelec.py tests/test_gpu_edge_cases.py tests/test_zz_no_alloc.py -m gpu -q -x > gpurun_out/r2as_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2as_pytest.log
tail -3 gpurun_out/r2as_pytest.log
