#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_host.py tests/test_gpu_parity.py -m gpu -x -q -k "simulator or fused" > gpurun_out/b2_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/b2_tests.log
tail -5 gpurun_out/b2_tests.log
python bench.py --steps 20 --warmup 3 > gpurun_out/b2_bench.json 2> gpurun_out/b2_bench.err; echo "bench rc=$?"
tail -3 gpurun_out/b2_bench.err
python - <<'P'
import json
for l in open('gpurun_out/b2_bench.json'):
    if l.startswith('{'):
        j=json.loads(l)
        print('ms/step %.3f value %.4g e2e %.4g blocking %.4g launches %d'%(j['ms_per_step'],j['value'],j['e2e']['value'],j['e2e']['blocking_value'],j['gpu_launches']))
        print({k:round(v*j['ms_per_step'],3) for k,v in j['roofline']['kernel_share_of_step'].items()})
        for k,v in j.get('configs',{}).items(): print(k, {a:(round(b,4) if isinstance(b,float) else b) for a,b in v.items()})
        print(j.get('cpu_baseline'))
P
ncu --set full --clock-control none --import-source on -k regex:"k_find_neighbors_list|k_pressure_force" -s 12 -c 4 -f -o gpurun_out/r02_src python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-configs > gpurun_out/b2_ncu.log 2>&1; echo "ncu rc=$?"
ls -la gpurun_out/
