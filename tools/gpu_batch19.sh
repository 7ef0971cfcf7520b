#!/bin/bash
# lanes-per-particle gather kernels at 8 M particles (default there: 1 lane)
mkdir -p gpurun_out
for l in 1 4 8; do
  SPH_LANES=$l timeout 200 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-configs > gpurun_out/b19_l$l.json 2> gpurun_out/b19_l$l.err
  python - $l <<'PY'
import json,sys
v=sys.argv[1]
try:
    j=json.loads(open('gpurun_out/b19_l%s.json'%v).read().strip().splitlines()[-1])
    sh=j['roofline']['kernel_share_of_step']; ms=j['ms_per_step']
    print('lanes',v,'ms/step %.4f'%ms, {k:round(x*ms,3) for k,x in sh.items()})
except Exception as e: print(v,'failed',e)
PY
done
