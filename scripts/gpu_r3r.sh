mkdir -p gpurun_out
export FK_NO_P2P=1
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 --full-solve 0 > gpurun_out/r3r_bench2_nccl.json 2> gpurun_out/r3r_bench2_nccl.err
echo "nccl exit $?"; cut -c1-200 gpurun_out/r3r_bench2_nccl.json
unset FK_NO_P2P
timeout 600 python bench.py --n 150 --no-cpu-baseline --full-solve 0 > gpurun_out/r3r_bench_n150.json 2> gpurun_out/r3r_bench_n150.err
echo "n150 exit $?"; cut -c1-200 gpurun_out/r3r_bench_n150.json; tail -n 2 gpurun_out/r3r_bench_n150.err
