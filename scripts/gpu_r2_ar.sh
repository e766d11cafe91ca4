#!/bin/bash
# final state: whole GPU suite (1 GPU) + smoke()
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/r2ar_pytest_gpu.log 2>&1; echo "pytest -m gpu rc=$?" | tee -a gpurun_out/r2ar_pytest_gpu.log
tail -4 gpurun_out/r2ar_pytest_gpu.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
