#!/bin/bash
# round 2, call k (1 GPU): per-kernel durations inside the bench step with the caches left as the step leaves them
mkdir -p gpurun_out
python scripts/asm_bench.py 119 10 > gpurun_out/r2k_asm_bench.jsonl 2> gpurun_out/r2k_asm_bench.err; echo "asm_bench rc=$?"
cut -c1-200 gpurun_out/r2k_asm_bench.jsonl
timeout 600 python bench.py --steps 3 --warmup 3 --full-solve 0 --no-cpu-baseline --trace 0 --pcg-iters 20 > gpurun_out/r2k_plain.json 2> gpurun_out/r2k_plain.err &&
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --cache-control none -k regex:"geo_rescale|picard_element|material_update|allreduce|fold_partials|max_array|assemble_kernel" --csv --log-file gpurun_out/r2k_launches.csv python bench.py --steps 3 --warmup 3 --full-solve 0 --no-cpu-baseline --trace 0 --pcg-iters 20 > gpurun_out/r2k_ncu.log 2>&1; echo "ncu rc=$?"
tail -30 gpurun_out/r2k_launches.csv | cut -c1-220
