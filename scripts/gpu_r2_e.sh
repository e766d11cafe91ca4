#!/bin/bash
# round 2, call e (2 GPUs): rescale assembly + LL mailbox: GPU suite, 1-GPU bench, 2-GPU benches with the timeline
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x --durations=8 > gpurun_out/r2e_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2e_pytest.log
tail -25 gpurun_out/r2e_pytest.log
timeout 600 python bench.py --steps 10 --warmup 3 --full-solve 0 --no-cpu-baseline > gpurun_out/r2e_1gpu.json 2> gpurun_out/r2e_1gpu.err; echo "1gpu rc=$?"
timeout 600 python bench.py --steps 10 --warmup 3 --full-solve 0 --no-cpu-baseline --tune 10=0 > gpurun_out/r2e_1gpu_tile.json 2> gpurun_out/r2e_1gpu_tile.err; echo "1gpu tile rc=$?"
timeout 600 python bench.py --steps 10 --warmup 3 --full-solve 0 --no-cpu-baseline --n 150 > gpurun_out/r2e_1gpu_n150.json 2> gpurun_out/r2e_1gpu_n150.err; echo "1gpu n150 rc=$?"
run() { local name=$1; shift
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 --full-solve 0 "$@" > gpurun_out/r2e_$name.json 2> gpurun_out/r2e_$name.err; echo "$name rc=$?"
}
run rec0_fused --recurrence 0
run rec0_split --recurrence 0 --tune 5=1
run rec1 --recurrence 1
