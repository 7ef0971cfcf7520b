#!/bin/bash
# Round-2 evidence run on ONE B200: tests, the default bench line, the reference arm, launch list, full ncu capture.
# Outputs under gpurun_out/ (copied to profiles/ by hand after reading).
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r02_gputests.log 2>&1; echo "pytest exit $?" >> gpurun_out/r02_gputests.log
tail -3 gpurun_out/r02_gputests.log
python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/r02_bench_reference.json 2> gpurun_out/r02_bench_reference.err; echo "reference rc=$?"
python bench.py --steps 20 --warmup 3 > gpurun_out/r02_bench_n1.json 2> gpurun_out/r02_bench_n1.err; echo "bench rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-configs > gpurun_out/r02_ncu_list.log 2>&1; echo "ncu list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:"k_find_neighbors_list|k_pressure_force|k_forces|k_predict_density|k_rank_reorder|k_hash_count|k_scatter|k_soa" -s 39 -c 13 -f -o gpurun_out/r02_full python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-configs > gpurun_out/r02_ncu_full.log 2>&1; echo "ncu full rc=$?"
ncu -i gpurun_out/r02_full.ncu-rep --page raw --csv > gpurun_out/r02_ncu_raw.csv 2>/dev/null
ls -la gpurun_out/
