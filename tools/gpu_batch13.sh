#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/b13_tests.log 2>&1; echo "pytest exit $?" >> gpurun_out/b13_tests.log; tail -3 gpurun_out/b13_tests.log
one() { timeout 150 python bench.py --workload $1 --steps $3 --warmup 10 --no-cpu-baseline --no-configs 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        j=json.loads(l); k=j['roofline']['kernel_share_of_step']; print('$2 $1 ms/step', round(j['ms_per_step'],4), '| eager:', {a:round(b*j['ms_per_step'],4) for a,b in k.items()})
"; }
one config_a base 300
for v in cp4 cp5 cp6; do SPH_LIB=$PWD/smoothed-particle-hydrodynamics_b200/build_exp/libsph_$v.so one config_a $v 300; done
one fluid_8m base 20
