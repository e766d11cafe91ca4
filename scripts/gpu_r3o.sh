mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_multi.py tests/test_gpu.py -m gpu -x -q -k "multi or pcg" > gpurun_out/r3o_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r3o_pytest.log
tail -n 4 gpurun_out/r3o_pytest.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 --full-solve 0 > gpurun_out/r3o_bench2_p2p.json 2> gpurun_out/r3o_bench2_p2p.err
echo "bench2 exit $?"; cut -c1-1300 gpurun_out/r3o_bench2_p2p.json; tail -n 3 gpurun_out/r3o_bench2_p2p.err
