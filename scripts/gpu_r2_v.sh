#!/bin/bash
# does the kind of L2 flush matter?  zeros (cudaMemset-like fill) vs a copy of random bytes
mkdir -p gpurun_out
FK_WINDOWS=65536 FK_FLUSH=random python scripts/asm_bench.py 119 10 short > gpurun_out/r2v_asm_bench_random_flush.jsonl 2> gpurun_out/r2v.err; echo rc=$?
cut -c1-200 gpurun_out/r2v_asm_bench_random_flush.jsonl; tail -3 gpurun_out/r2v.err
FK_WINDOWS=65536 python scripts/asm_bench.py 119 10 short > gpurun_out/r2v_asm_bench_zero_flush.jsonl 2>> gpurun_out/r2v.err; echo rc=$?
cut -c1-200 gpurun_out/r2v_asm_bench_zero_flush.jsonl
