# usage: bash scripts/gpu_r3q.sh N   (N GPUs, weak-scaling bench over NVLink peer memory)
N=$1
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 --full-solve 0 > gpurun_out/r3q_bench${N}.json 2> gpurun_out/r3q_bench${N}.err
echo "bench$N exit $?"; cut -c1-1300 gpurun_out/r3q_bench${N}.json; tail -n 3 gpurun_out/r3q_bench${N}.err
