mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r1p_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r1p_pytest.log
tail -n 5 gpurun_out/r1p_pytest.log
timeout 600 python bench.py --steps 5 --warmup 3 --full-solve 0 --no-cpu-baseline > gpurun_out/r1p_bench1.json 2> gpurun_out/r1p_bench1.err; echo "bench1 exit $?"
grep -o "\"value\": [0-9.]*\|\"pcg_ms\": [0-9.]*" gpurun_out/r1p_bench1.json
for mode in p2p nccl; do
  if [ $mode = nccl ]; then export FK_NO_P2P=1; else unset FK_NO_P2P; fi
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 --full-solve 0 > gpurun_out/r1p_bench2_$mode.json 2> gpurun_out/r1p_bench2_$mode.err
  echo "bench $mode exit $?"; grep -o "\"value\": [0-9.]*\|\"pcg_ms\": [0-9.]*" gpurun_out/r1p_bench2_$mode.json; tail -n 3 gpurun_out/r1p_bench2_$mode.err | cut -c1-300
done
