#!/bin/bash
# round 2, call r (2 GPUs): heat workload on 2 GPUs (configs[2]), tightened scalarPotential drop-in test
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu.py -m gpu -q -s -k "scalar_potential" > gpurun_out/r2r_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2r_pytest.log
grep -v "^$" gpurun_out/r2r_pytest.log | tail -12
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --workload heat --cells 150 --steps 20 --warmup 3 > gpurun_out/r2r_heat_2gpu.json 2> gpurun_out/r2r_heat_2gpu.err; echo "heat2 rc=$?"
grep "^{" gpurun_out/r2r_heat_2gpu.json | cut -c1-1200; tail -2 gpurun_out/r2r_heat_2gpu.err
