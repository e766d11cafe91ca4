mkdir -p gpurun_out
echo NEW; timeout 300 python scripts/dbg_dropin.py 2>&1 | tail -n 16
echo OLD; timeout 300 python scripts/dbg_dropin.py scripts/_tmp/libfemeko_old.so 2>&1 | tail -n 16
