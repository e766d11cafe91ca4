#!/bin/bash
mkdir -p gpurun_out
timeout 80 python -m pytest tests/test_gpu.py tests/test_zz_no_alloc.py -m gpu -q -x -k "golden or assembly_family or pcg_against or noalloc" > gpurun_out/r2at_pytest.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2at_pytest.log
tail -3 gpurun_out/r2at_pytest.log
