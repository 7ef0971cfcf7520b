#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/b3_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/b3_tests.log
tail -15 gpurun_out/b3_tests.log
python bench.py --steps 20 --warmup 3 --no-configs --no-cpu-baseline > gpurun_out/b3_bench.json 2> gpurun_out/b3_bench.err; echo "bench rc=$?"
for t in p2p nccl; do
  SPH_SLAB_TRANSPORT=$t python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/b3_bench_n2_$t.json 2> gpurun_out/b3_bench_n2_$t.err; echo "n2 $t rc=$?"
  tail -3 gpurun_out/b3_bench_n2_$t.err
done
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/b3_bench*.json')):
    for l in open(f):
        if l.startswith('{'):
            j=json.loads(l)
            print(f,'ms/step %.3f value %.4g e2e %.4g blocking %.4g launches %d'%(j['ms_per_step'],j['value'],j['e2e']['value'],j['e2e']['blocking_value'],j['gpu_launches']))
            print({k:round(v*j['ms_per_step'],3) for k,v in j['roofline']['kernel_share_of_step'].items()})
            print(j.get('slab_parity'), j.get('strong_16m'))
P
