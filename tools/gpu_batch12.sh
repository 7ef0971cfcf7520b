#!/bin/bash
mkdir -p gpurun_out
T="timeout 150"
$T python -m pytest tests/test_gpu_parity.py tests/test_host.py tests/test_zz_no_alloc.py tests/test_slab_gpu.py -m gpu -x -q 2>&1 | tail -2
one() { $T python bench.py --workload $1 --steps $3 --warmup 10 --no-cpu-baseline --no-configs 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        j=json.loads(l); k=j['roofline']['kernel_share_of_step']; print('$2 $1 ms/step %.4f | eager:', {a:round(b*j['ms_per_step'],4) for a,b in k.items()})
"; echo "rc=$? $2 $1"; }
one config_a base 300
one config_b base 200
one fluid_8m base 20
for v in cp4 cp5 cp6 t64; do SPH_LIB=$PWD/smoothed-particle-hydrodynamics_b200/build_exp/libsph_$v.so one config_a $v 300; done
for v in f4 f5 f7; do SPH_LIB=$PWD/smoothed-particle-hydrodynamics_b200/build_exp/libsph_$v.so one fluid_8m $v 20; done
timeout 400 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/b12_bench.json 2> gpurun_out/b12_bench.err; echo "bench rc=$?"
python - <<'P'
import json
for l in open('gpurun_out/b12_bench.json'):
    if l.startswith('{'):
        j=json.loads(l)
        print('ms/step %.3f value %.4g e2e %.4g blocking %.4g pfrac %s'%(j['ms_per_step'],j['value'],j['e2e']['value'],j['e2e']['blocking_value'], j['config'].get('pressure_nonzero_frac')))
        for k,v in j.get('configs',{}).items(): print(k, {a:(round(b,4) if isinstance(b,float) else b) for a,b in v.items() if a in ('ms_per_step','value','pressure_nonzero_frac','later_with_pressure')})
P
