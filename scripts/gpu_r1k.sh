mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/r1m_topo.txt 2>&1
timeout 600 python -m pytest tests/test_gpu_multi.py -m gpu -x -q > gpurun_out/r1m_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r1m_pytest.log
tail -n 40 gpurun_out/r1m_pytest.log
for mode in p2p nccl; do
  if [ $mode = nccl ]; then export FK_NO_P2P=1; else unset FK_NO_P2P; fi
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 --full-solve 0 > gpurun_out/r1m_bench2_$mode.json 2> gpurun_out/r1m_bench2_$mode.err
  echo "bench $mode exit $?"; cat gpurun_out/r1m_bench2_$mode.json | cut -c1-1500; tail -n 5 gpurun_out/r1m_bench2_$mode.err
done
