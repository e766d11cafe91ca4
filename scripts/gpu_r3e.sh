mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r3e_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r3e_pytest.log
tail -n 3 gpurun_out/r3e_pytest.log
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/r3e_bench.json 2> gpurun_out/r3e_bench.err; echo "bench exit $?" >> gpurun_out/r3e_bench.err
cut -c1-1500 gpurun_out/r3e_bench.json; tail -n 3 gpurun_out/r3e_bench.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 --full-solve 0 > gpurun_out/r3e_bench2_p2p.json 2> gpurun_out/r3e_bench2_p2p.err
echo "bench2 exit $?"; cut -c1-1700 gpurun_out/r3e_bench2_p2p.json; tail -n 5 gpurun_out/r3e_bench2_p2p.err
