#!/bin/bash
mkdir -p gpurun_out
python tools/diag_worm.py 2>&1 | tail -7
python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -3
for v in 0 1; do for wl in config_a config_b; do SPH_PDL=$v python bench.py --workload $wl --steps 400 --warmup 10 --no-cpu-baseline --no-configs 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        j=json.loads(l); print('PDL=$v $wl ms/step %.4f'%j['ms_per_step'])
"; done; done
python -m pytest tests -m gpu -x -q > gpurun_out/b7_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/b7_tests.log
tail -5 gpurun_out/b7_tests.log
