#!/bin/bash
# round 2, call d (2 GPUs): device-side timeline of the PCG iteration on 1 and 2 GPUs (where do the 20 us of the 1 -> 2 cliff go?)
mkdir -p gpurun_out
timeout 600 python bench.py --steps 5 --warmup 3 --full-solve 0 --no-cpu-baseline > gpurun_out/r2d_1gpu_rec0.json 2> gpurun_out/r2d_1gpu_rec0.err; echo "1gpu rec0 rc=$?"
timeout 600 python bench.py --steps 5 --warmup 3 --full-solve 0 --no-cpu-baseline --recurrence 1 > gpurun_out/r2d_1gpu_rec1.json 2> gpurun_out/r2d_1gpu_rec1.err; echo "1gpu rec1 rc=$?"
run() { local name=$1; shift
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 --full-solve 0 "$@" > gpurun_out/r2d_$name.json 2> gpurun_out/r2d_$name.err; echo "$name rc=$?"
}
run rec0_fused --recurrence 0
run rec0_split --recurrence 0 --tune 5=1
run rec1 --recurrence 1
timeout 300 python -m pytest tests/test_gpu_multi.py -m gpu -q -k "0-1" 2>&1 | tail -3
