#!/bin/bash
# first GPU batch of round 2: parity of the new kernels, small-scene timings per variant
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/b1_parity.log 2>&1; echo "parity rc=$?" >> gpurun_out/b1_parity.log
tail -5 gpurun_out/b1_parity.log
for wl in config_a config_b; do
  for v in "SPH_GRAPH=0 SPH_LANES=1 SPH_NEIGHBOR_PATH=list" "SPH_GRAPH=1 SPH_LANES=1 SPH_NEIGHBOR_PATH=list" \
           "SPH_GRAPH=1 SPH_LANES=4 SPH_NEIGHBOR_PATH=list" "SPH_GRAPH=1 SPH_LANES=8 SPH_NEIGHBOR_PATH=list" \
           "SPH_GRAPH=1 SPH_LANES=16 SPH_NEIGHBOR_PATH=list" "SPH_GRAPH=1 SPH_LANES=8 SPH_NEIGHBOR_PATH=coop" \
           "SPH_GRAPH=1 SPH_LANES=4 SPH_NEIGHBOR_PATH=coop" "SPH_GRAPH=0 SPH_LANES=8 SPH_NEIGHBOR_PATH=coop"; do
    echo "== $wl $v" >> gpurun_out/b1_small.log
    env $v python bench.py --workload $wl --steps 200 --warmup 10 --no-cpu-baseline 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        j=json.loads(l); print('ms/step %.4f value %.4g e2e %.4g blocking %.4g'%(j['ms_per_step'],j['value'],j['e2e']['value'],j['e2e']['blocking_value'])); print({k:round(v*j['ms_per_step'],4) for k,v in j['roofline']['kernel_share_of_step'].items()})
    else: print(l.rstrip())
" >> gpurun_out/b1_small.log
  done
done
cat gpurun_out/b1_small.log
