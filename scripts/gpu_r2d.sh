mkdir -p gpurun_out
PROF="python bench.py --steps 1 --warmup 3 --pcg-iters 4 --presolve 0 --full-solve 0 --no-cpu-baseline"
timeout 300 $PROF > gpurun_out/r2d_prof_plain.log 2>&1 && timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/r2d_launches.csv $PROF > gpurun_out/r2d_ncu1.log 2>&1
grep assemble_kernel gpurun_out/r2d_launches.csv | cut -d, -f5,14-16 | head -8
timeout 600 python -m pytest tests/test_gpu_multi.py -m gpu -x -q > gpurun_out/r2d_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r2d_pytest.log
tail -n 5 gpurun_out/r2d_pytest.log
for mode in p2p nccl; do
  if [ $mode = nccl ]; then export FK_NO_P2P=1; else unset FK_NO_P2P; fi
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 --full-solve 0 > gpurun_out/r2d_bench2_$mode.json 2> gpurun_out/r2d_bench2_$mode.err
  echo "bench $mode exit $?"; grep -o "\"value\": [0-9.]*\|\"pcg_ms\": [0-9.]*\|\"picard_assembly_ms\": [0-9.]*" gpurun_out/r2d_bench2_$mode.json; tail -n 3 gpurun_out/r2d_bench2_$mode.err | cut -c1-300
done
