#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_host.py tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -2
one() { python bench.py --workload $1 --steps 400 --warmup 20 --no-cpu-baseline --no-configs 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        j=json.loads(l); k=j['roofline']['kernel_share_of_step']; print('$2 $1 graph ms/step %.4f | eager: NS %.4f PF %.4f PFI %.4f forces %.4f PD %.4f'%(j['ms_per_step'], k['find_neighbors+density']*j['ms_per_step'], k['pressure_force+predict']*j['ms_per_step'], k['pressure_force+integrate']*j['ms_per_step'], k['forces+predict0']*j['ms_per_step'], k['predict_density+correct_pressure']*j['ms_per_step']))
"; }
one config_a base
for v in cp4 cp5 cp6 t64; do SPH_LIB=$PWD/smoothed-particle-hydrodynamics_b200/build_exp/libsph_$v.so one config_a $v; done
SPH_LANES=8 one config_a lanes8
python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/b11_bench.json 2> gpurun_out/b11_bench.err; echo "bench rc=$?"
python - <<'P'
import json
for l in open('gpurun_out/b11_bench.json'):
    if l.startswith('{'):
        j=json.loads(l)
        print('ms/step %.3f value %.4g e2e %.4g blocking %.4g pfrac %s'%(j['ms_per_step'],j['value'],j['e2e']['value'],j['e2e']['blocking_value'], j['config'].get('pressure_nonzero_frac')))
        for k,v in j.get('configs',{}).items(): print(k, {a:(round(b,4) if isinstance(b,float) else b) for a,b in v.items() if a in ('ms_per_step','value','pressure_nonzero_frac','after_2000_steps')})
P
