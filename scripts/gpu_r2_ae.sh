#!/bin/bash
# DRAM bytes of the sweep's kernels inside the bench step (row-sum diagonal, side-by-side sweep), caches left as the step leaves them
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --full-solve 0 --trace 0"
$CMD > gpurun_out/r2ae_plain.json 2> gpurun_out/r2ae_plain.err &&
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct --clock-control none --cache-control none --graph-profiling node -k regex:"geo_rescale|diag_from_rowsum|picard_element|material_update|max_array|geo_scale" -c 80 --csv --log-file gpurun_out/r2ae_sweep_kernels.csv $CMD > gpurun_out/r2ae_ncu.log 2>&1; echo "ncu rc=$?"
tail -3 gpurun_out/r2ae_ncu.log; wc -l gpurun_out/r2ae_sweep_kernels.csv
