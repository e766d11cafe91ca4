mkdir -p gpurun_out
for T in "5=0" "6=0" "6=1" "6=2"; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 --full-solve 0 --presolve 0 --tune "$T" > gpurun_out/r3i_bench2_$T.json 2> gpurun_out/r3i_bench2_$T.err
  echo "tune $T exit $?"; python -c "
import json,sys
d=json.loads(open('gpurun_out/r3i_bench2_$T.json').read().strip().splitlines()[-1])
print(d['value'], d['components'], d['gpu_launches'])"
done
