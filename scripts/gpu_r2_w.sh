#!/bin/bash
# which regime is the sweep in?  per-launch clock, DRAM bytes and L2 hit rate of the sweep kernels, caches left alone
mkdir -p gpurun_out
FK_WINDOWS=65536 python scripts/asm_bench.py 119 3 short > gpurun_out/r2w_plain.log 2>&1 &&
FK_WINDOWS=65536 ncu --metrics gpu__time_duration.sum,sm__cycles_elapsed.avg.per_second,dram__cycles_elapsed.avg.per_second,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,lts__t_sectors_op_write.sum,lts__t_sectors_op_read.sum --clock-control none --cache-control none -k regex:"geo_rescale|picard_element" --csv --log-file gpurun_out/r2w_regimes.csv python scripts/asm_bench.py 119 3 short > gpurun_out/r2w_ncu.log 2>&1; echo "ncu rc=$?"
tail -2 gpurun_out/r2w_ncu.log
