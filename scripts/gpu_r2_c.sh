#!/bin/bash
# round 2, call c (1 GPU): L2 policy experiments for both recurrences
mkdir -p gpurun_out
run() { local name=$1; shift
  FK_DEBUG=1 timeout 600 python bench.py --steps 10 --warmup 3 --full-solve 0 --no-cpu-baseline "$@" > gpurun_out/r2c_$name.json 2> gpurun_out/r2c_$name.err; echo "$name rc=$?"; }
run rec1 --recurrence 1
run rec1_l2normal --recurrence 1 --tune 8=1
run rec1_v4 --recurrence 1 --tune 9=4
run rec1_v4_normal --recurrence 1 --tune 9=4,8=1
run rec1_v6 --recurrence 1 --tune 9=6
run rec0 --recurrence 0
run rec0_l2normal --recurrence 0 --tune 8=1
run n150_rec0 --recurrence 0 --n 150
run n150_rec0_normal --recurrence 0 --n 150 --tune 8=1
run n150_rec1 --recurrence 1 --n 150
run n150_rec1_normal --recurrence 1 --n 150 --tune 8=1
timeout 300 python -m pytest tests/test_gpu_round2.py -m gpu -q -k "dirichlet" > gpurun_out/r2c_pytest.log 2>&1; tail -3 gpurun_out/r2c_pytest.log
grep -h "fk: L2" gpurun_out/r2c_rec1.err | head -3
