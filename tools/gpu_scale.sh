#!/bin/bash
# usage: tools/gpu_scale.sh N   — bench.py on N GPUs (weak fluid_8m per GPU + strong fluid_16m + slab parity self-check)
N=$1
mkdir -p gpurun_out
SPH_BENCH_MARKS=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/r02_bench_n$N.json 2> gpurun_out/r02_bench_n$N.err; echo "n$N rc=$?"
grep -v "^\*\*\*\|OMP_NUM" gpurun_out/r02_bench_n$N.err | tail -12
python - <<P
import json
for l in open('gpurun_out/r02_bench_n$N.json'):
    if l.startswith('{'):
        j=json.loads(l)
        print('N=%d ms/step %.3f value %.4g e2e %.4g blocking %.4g launches %d'%(j['n_gpus'],j['ms_per_step'],j['value'],j['e2e']['value'],j['e2e']['blocking_value'],j['gpu_launches']))
        print({k:round(v*j['ms_per_step'],3) for k,v in j['roofline']['kernel_share_of_step'].items()})
        print(j.get('slab_parity')); print(j.get('strong_16m')); print(j.get('cpu_binding'), j['clocks'])
P
