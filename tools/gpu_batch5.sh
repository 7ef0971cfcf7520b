#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_host.py tests/test_slab_gpu.py -m gpu -x -q > gpurun_out/b5_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/b5_tests.log
tail -4 gpurun_out/b5_tests.log
python bench.py --steps 20 --warmup 3 --no-configs --no-cpu-baseline > gpurun_out/b5_bench.json 2> gpurun_out/b5_bench.err; echo "bench rc=$?"
SPH_BENCH_MARKS=1 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 --no-configs > gpurun_out/b5_bench_n2.json 2> gpurun_out/b5_bench_n2.err; echo "n2 rc=$?"
grep marks gpurun_out/b5_bench_n2.err
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/b5_bench*.json')):
    for l in open(f):
        if l.startswith('{'):
            j=json.loads(l)
            print(f,'ms/step %.3f value %.4g e2e %.4g blocking %.4g launches %d'%(j['ms_per_step'],j['value'],j['e2e']['value'],j['e2e']['blocking_value'],j['gpu_launches']))
            print({k:round(v*j['ms_per_step'],3) for k,v in j['roofline']['kernel_share_of_step'].items()})
            print(j.get('slab_parity'))
P
