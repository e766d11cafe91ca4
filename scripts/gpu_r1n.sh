mkdir -p gpurun_out
export FK_COMM_TRACE=1
for mode in p2p nccl; do
  if [ $mode = nccl ]; then export FK_NO_P2P=1; else unset FK_NO_P2P; fi
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 2 --warmup 3 --full-solve 0 --presolve 0 > gpurun_out/r1n_bench2_$mode.json 2> gpurun_out/r1n_bench2_$mode.err
  echo "bench $mode exit $?"; grep "comm trace" gpurun_out/r1n_bench2_$mode.err | tail -n 4
done
