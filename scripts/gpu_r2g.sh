mkdir -p gpurun_out
N=$1
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 --full-solve 0 > gpurun_out/r2g_bench$N.json 2> gpurun_out/r2g_bench$N.err
echo "bench $N exit $?"; grep -o "\"value\": [0-9.]*\|\"pcg_ms\": [0-9.]*\|\"picard_assembly_ms\": [0-9.]*\|\"parallelism\": \"[^\"]*\"" gpurun_out/r2g_bench$N.json; grep -v "^\*\*\*\|OMP_NUM" gpurun_out/r2g_bench$N.err | tail -n 8 | cut -c1-300
