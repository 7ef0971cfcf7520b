#!/bin/bash
# neighbour-search variants: per-entry base (e0), shared-memory cell bases at 9 / 10 CTAs per SM
mkdir -p gpurun_out
D=smoothed-particle-hydrodynamics_b200/build_exp
for v in ${VARS:-e0 cbs10}; do
  SPH_LIB=$D/libsph_$v.so timeout 200 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-configs > gpurun_out/b18_$v.json 2> gpurun_out/b18_$v.err
  python - $v <<'PY'
import json,sys
v=sys.argv[1]
try:
    j=json.loads(open('gpurun_out/b18_%s.json'%v).read().strip().splitlines()[-1])
    print(v, 'ms/step %.4f'%j['ms_per_step'], 'NS ms %.4f'%j['roofline']['avg_launch_ms'])
except Exception as e: print(v,'failed',e)
PY
done
true
