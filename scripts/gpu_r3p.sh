mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r3p_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r3p_pytest.log
tail -n 3 gpurun_out/r3p_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r3p_smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/r3p_smoke.log
tail -n 2 gpurun_out/r3p_smoke.log
timeout 900 python bench.py > gpurun_out/r3p_bench.json 2> gpurun_out/r3p_bench.err; echo "bench exit $?" >> gpurun_out/r3p_bench.err
cat gpurun_out/r3p_bench.json | cut -c1-3000; tail -n 3 gpurun_out/r3p_bench.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r3p_ref.json 2> gpurun_out/r3p_ref.err
cut -c1-600 gpurun_out/r3p_ref.json
PROF="python bench.py --steps 1 --warmup 3 --pcg-iters 4 --presolve 0 --full-solve 0 --no-cpu-baseline"
timeout 300 $PROF > gpurun_out/r3p_prof_plain.log 2>&1 && timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/r3p_launches.csv $PROF > gpurun_out/r3p_ncu1.log 2>&1
timeout 300 $PROF > gpurun_out/r3p_prof_plain2.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:'assemble_kernel|material_update|spmv_sell|pcg_update|pcg_direction' -s 10 -c 10 -f -o gpurun_out/r3p_prof $PROF > gpurun_out/r3p_ncu2.log 2>&1
ls -la gpurun_out | tail -n 12
