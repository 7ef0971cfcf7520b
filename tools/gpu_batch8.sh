#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "random_gas or worm_fused or config_a_fused or synthetic" 2>&1 | tail -2
for wl in config_a config_b dambreak_1m fluid_8m; do SPH_GRAPH=0 python bench.py --workload $wl --steps 50 --warmup 5 --no-cpu-baseline --no-configs 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        j=json.loads(l); print('$wl ms/step %.4f e2e %.4g blocking %.4g'%(j['ms_per_step'], j['e2e']['value'], j['e2e']['blocking_value'])); print({k:round(v*j['ms_per_step'],4) for k,v in j['roofline']['kernel_share_of_step'].items()})
"; done
python bench.py --steps 20 --warmup 3 > gpurun_out/b8_bench.json 2> gpurun_out/b8_bench.err; echo "bench rc=$?"
python - <<'P'
import json
for l in open('gpurun_out/b8_bench.json'):
    if l.startswith('{'):
        j=json.loads(l)
        print('ms/step %.3f value %.4g e2e %.4g blocking %.4g launches %d'%(j['ms_per_step'],j['value'],j['e2e']['value'],j['e2e']['blocking_value'],j['gpu_launches']))
        for k,v in j.get('configs',{}).items(): print(k, {a:(round(b,4) if isinstance(b,float) else b) for a,b in v.items() if a in ('ms_per_step','value','e2e_value','e2e_blocking_value','x_cpu','x_cpu_e2e','cpu_value')})
P
