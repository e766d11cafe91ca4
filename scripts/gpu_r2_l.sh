#!/bin/bash
# round 2, call l (1 GPU): does the sweep slow down behind a PCG solve because of its persisting L2 lines?  + 2-D P1 tests
mkdir -p gpurun_out
FK_WINDOWS=65536 python scripts/asm_bench.py 119 10 > gpurun_out/r2l_asm_bench.jsonl 2> gpurun_out/r2l_asm_bench.err; echo "asm_bench rc=$?"
cut -c1-200 gpurun_out/r2l_asm_bench.jsonl; tail -3 gpurun_out/r2l_asm_bench.err
