#!/bin/bash
mkdir -p gpurun_out
D=smoothed-particle-hydrodynamics_b200/build_exp
for v in ${VARS}; do
  SPH_LIB=$D/libsph_$v.so timeout 200 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-configs > gpurun_out/b20_$v.json 2> gpurun_out/b20_$v.err
  python - $v <<'PY'
import json,sys
v=sys.argv[1]
try:
    j=json.loads(open('gpurun_out/b20_%s.json'%v).read().strip().splitlines()[-1])
    print(v, 'ms/step %.4f'%j['ms_per_step'], 'NS ms %.4f'%j['roofline']['avg_launch_ms'])
except Exception as e: print(v,'failed',e)
PY
done
[ -n "$PARITY" ] && { SPH_LIB=$D/libsph_$PARITY.so timeout 300 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py -m gpu -x -q -k "default or histonly or fullsize or 8m" > gpurun_out/b20_parity.log 2>&1; tail -2 gpurun_out/b20_parity.log; }
