#!/bin/bash
# round 2, call m (1 GPU): is the sweep slow behind a host sync because the GPU was idle?  + 2-D P1 tests on the GPU
mkdir -p gpurun_out
FK_WINDOWS=65536 python scripts/asm_bench.py 119 10 > gpurun_out/r2m_asm_bench.jsonl 2> gpurun_out/r2m_asm_bench.err; echo "asm_bench rc=$?"
cut -c1-200 gpurun_out/r2m_asm_bench.jsonl; tail -3 gpurun_out/r2m_asm_bench.err
timeout 600 python -m pytest tests/test_gpu_2d.py tests/test_gpu_rescale.py -m gpu -q -x > gpurun_out/r2m_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2m_pytest.log
tail -12 gpurun_out/r2m_pytest.log
