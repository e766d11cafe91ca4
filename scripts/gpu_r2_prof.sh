#!/bin/bash
# round 2 profiles (1 GPU): launch list of the bench command, steady-state per-kernel DRAM bytes inside the replayed PCG graph
# and the sweep (caches left as the step leaves them), one full capture of the top kernels
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --full-solve 0 --trace 0"
$CMD > gpurun_out/r2p_plain.json 2> gpurun_out/r2p_plain.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --graph-profiling node --csv --log-file gpurun_out/r2p_launches.csv $CMD > gpurun_out/r2p_ncu1.log 2>&1; echo "ncu launches rc=$?"
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct --clock-control none --cache-control none --graph-profiling node -k regex:"spmv_sell|pcg_update|pcg_direction|geo_rescale|picard_element|material_update" -s 900 -c 340 --csv --log-file gpurun_out/r2p_steady.csv $CMD > gpurun_out/r2p_ncu2.log 2>&1; echo "ncu steady rc=$?"
ncu --set full --clock-control none --import-source on --graph-profiling node -k regex:"spmv_sell_kernel|pcg_update_kernel|pcg_direction_kernel|geo_rescale_kernel|picard_element_kernel" -s 1200 -c 10 -o gpurun_out/r2p_full $CMD > gpurun_out/r2p_ncu3.log 2>&1; echo "ncu full rc=$?"
ls -la gpurun_out/r2p_*; tail -3 gpurun_out/r2p_ncu1.log gpurun_out/r2p_ncu2.log gpurun_out/r2p_ncu3.log
