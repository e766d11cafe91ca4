#!/bin/bash
# round 2, call b (2 GPUs): partitioned parity in every mode + 2-GPU bench lines: Eigen fused / Eigen split / single-reduction
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_round2.py -m gpu -q -k "dirichlet" > gpurun_out/r2b_pytest_dir.log 2>&1; tail -3 gpurun_out/r2b_pytest_dir.log
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -q -s > gpurun_out/r2b_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2b_pytest.log
run() { # name, args...
  local name=$1; shift
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 --full-solve 0 "$@" > gpurun_out/r2b_$name.json 2> gpurun_out/r2b_$name.err; echo "$name rc=$?"
}
run rec0_fused --recurrence 0
run rec0_split --recurrence 0 --tune 5=1
run rec1 --recurrence 1
run rec1_l2normal --recurrence 1 --tune 8=1
tail -15 gpurun_out/r2b_pytest.log
