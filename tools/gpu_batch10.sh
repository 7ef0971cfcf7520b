#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_host.py -m gpu -x -q -k "worm_fused or config_a_fused or random_gas or bprime or several_slabs" 2>&1 | tail -2
one() { python bench.py --workload $1 --steps 30 --warmup 5 --no-cpu-baseline --no-configs 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        j=json.loads(l); k=j['roofline']['kernel_share_of_step']; print('$2 $1 ms/step %.4f NS %.4f PF %.4f PFI %.4f forces %.4f PD %.4f'%(j['ms_per_step'], k['find_neighbors+density']*j['ms_per_step'], k['pressure_force+predict']*j['ms_per_step'], k['pressure_force+integrate']*j['ms_per_step'], k['forces+predict0']*j['ms_per_step'], k['predict_density+correct_pressure']*j['ms_per_step']))
"; }
export SPH_GRAPH=0
for pct in 5 10 15 20 100; do for wl in config_b dambreak_1m fluid_8m; do SPH_HIST_PCT=$pct one $wl pct=$pct; done; done
for v in pf8 pf6 pfb8 f7 f5 ns12 ns8; do SPH_LIB=$PWD/smoothed-particle-hydrodynamics_b200/build_exp/libsph_$v.so one fluid_8m $v; done
