set -x
python tools/bench_configs.py > gpurun_out/configs.json 2> gpurun_out/configs.err || tail -5 gpurun_out/configs.err
python bench.py > gpurun_out/bench_final.log 2>&1; tail -1 gpurun_out/bench_final.log | cut -c1-600
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off -c 600 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --profile-region --no-e2e --no-cpu-baseline > gpurun_out/ncu_list.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:am_rollout_tsp -c 1 --profile-from-start off -f -o gpurun_out/r01_k2 python bench.py --steps 1 --warmup 3 --profile-region --no-e2e --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off -c 3000 --csv --log-file gpurun_out/e2e_launches.csv python bench.py --steps 1 --warmup 3 --profile-e2e --batch 131072 --no-cpu-baseline > gpurun_out/ncu_e2e.log 2>&1
ls -la gpurun_out | tail -12
