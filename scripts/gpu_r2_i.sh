#!/bin/bash
# round 2, call i (1 GPU): rescale plan windows beyond 4096 (do same-direction runs coalesce the s gathers?), heat workload
mkdir -p gpurun_out
python scripts/asm_bench.py 119 15 > gpurun_out/r2i_asm_bench.jsonl 2> gpurun_out/r2i_asm_bench.err; echo "asm_bench rc=$?"
cut -c1-200 gpurun_out/r2i_asm_bench.jsonl
timeout 900 python bench.py --workload heat --n 150 --steps 20 --warmup 3 > gpurun_out/r2i_heat_n150.json 2> gpurun_out/r2i_heat_n150.err; echo "heat rc=$?"
cut -c1-1500 gpurun_out/r2i_heat_n150.json; tail -3 gpurun_out/r2i_heat_n150.err
