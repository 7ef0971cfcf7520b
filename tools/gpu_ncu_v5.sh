#!/bin/bash
# ncu evidence of the final build: launch list + one full capture of a whole step (after the plain run exited 0)
mkdir -p gpurun_out
timeout 200 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-configs > gpurun_out/r02_v5_plain.json 2> gpurun_out/r02_v5_plain.err; echo "plain rc=$?"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_v5_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-configs > gpurun_out/r02_v5_ncu_list.log 2>&1; echo "ncu list rc=$?"
timeout 420 ncu --set full --clock-control none --import-source on -k regex:"k_find_neighbors_list|k_pressure_force|k_forces|k_predict_density|k_rank_reorder|k_hash_count|k_scatter|k_soa" -s 39 -c 13 -f -o gpurun_out/r02_v5_full python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-configs > gpurun_out/r02_v5_ncu_full.log 2>&1; echo "ncu full rc=$?"
timeout 120 ncu -i gpurun_out/r02_v5_full.ncu-rep --page raw --csv > gpurun_out/r02_v5_ncu_raw.csv 2>/dev/null
ls -la gpurun_out/ | grep v5
