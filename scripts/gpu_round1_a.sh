mkdir -p gpurun_out
(nproc; free -g | head -2; nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv; which julia gmsh; python -c "import gmsh" 2>&1 | tail -1) > gpurun_out/r1_sysinfo.txt 2>&1
python -m pytest tests -m gpu -x -q > gpurun_out/r1_pytest.log 2>&1; echo "pytest exit $?" >> gpurun_out/r1_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r1_smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/r1_smoke.log
python bench.py > gpurun_out/r1_bench.json 2> gpurun_out/r1_bench.err; echo "bench exit $?" >> gpurun_out/r1_bench.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r1_ref.json 2> gpurun_out/r1_ref.err
PROF="python bench.py --steps 1 --warmup 3 --pcg-iters 4 --presolve 0 --full-solve 0 --no-cpu-baseline"
$PROF > gpurun_out/r1_prof_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/r1_launches.csv $PROF > gpurun_out/r1_ncu1.log 2>&1
$PROF > gpurun_out/r1_prof_plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'assemble_kernel|spmv_kernel|pcg_update|pcg_direction' -s 13 -c 8 -o gpurun_out/r1_prof $PROF > gpurun_out/r1_ncu2.log 2>&1
ls -la gpurun_out
