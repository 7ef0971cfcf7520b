#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_slab_gpu.py tests/test_host.py -m gpu -x -q > gpurun_out/b9_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/b9_tests.log
tail -4 gpurun_out/b9_tests.log
for pct in 0 25 50 75 100; do for wl in config_b dambreak_1m fluid_8m; do SPH_HIST_PCT=$pct SPH_GRAPH=0 python bench.py --workload $wl --steps 30 --warmup 5 --no-cpu-baseline --no-configs 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        j=json.loads(l); k=j['roofline']['kernel_share_of_step']; print('pct=$pct $wl ms/step %.4f NS %.4f PF %.4f PFI %.4f forces %.4f PD %.4f'%(j['ms_per_step'], k['find_neighbors+density']*j['ms_per_step'], k['pressure_force+predict']*j['ms_per_step'], k['pressure_force+integrate']*j['ms_per_step'], k['forces+predict0']*j['ms_per_step'], k['predict_density+correct_pressure']*j['ms_per_step']))
"; done; done
