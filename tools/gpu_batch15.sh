#!/bin/bash
# histogram-only neighbour mode: parity, then the pressurised dam break with and without it; then the full GPU suite
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "histonly or default" > gpurun_out/b15_parity.log 2>&1; echo "parity rc=$?" >> gpurun_out/b15_parity.log
for h2 in 101 70 50 30; do
  SPH_HIST2_PCT=$h2 timeout 200 python tools/late_profile.py dambreak_1m 1200 > gpurun_out/b15_late_$h2.log 2>&1
done
SPH_HIST2_PCT=50 timeout 200 python bench.py --workload config_b --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/b15_cfgb_50.log 2>&1
SPH_HIST2_PCT=101 timeout 200 python bench.py --workload config_b --steps 50 --warmup 5 --no-cpu-baseline > gpurun_out/b15_cfgb_101.log 2>&1
tail -3 gpurun_out/b15_parity.log; grep -h "ms/step\|counters" gpurun_out/b15_late_*.log
