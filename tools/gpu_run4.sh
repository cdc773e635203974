# round-2 GPU pass 4: whole GPU suite, default bench, ncu evidence of the headline configuration (launch list, K2 full, e2e list)
set -x
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q -x --timeout 600 > gpurun_out/r02_pytest4.log 2>&1; echo "suite rc=$?"; tail -4 gpurun_out/r02_pytest4.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02_smoke.log 2>&1; echo "smoke rc=$?"
timeout 1200 python bench.py > gpurun_out/r02_bench3.log 2> gpurun_out/r02_bench3.err; echo "bench rc=$?"; tail -c 300 gpurun_out/r02_bench3.log; tail -3 gpurun_out/r02_bench3.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off -c 600 --csv --log-file gpurun_out/r02_launches_raw.csv python bench.py --steps 2 --warmup 3 --profile-region --no-e2e --no-cpu-baseline --no-side --no-parity > gpurun_out/ncu_list.log 2>&1; echo "list rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:am_rollout_tsp -c 1 --profile-from-start off -f -o gpurun_out/r02_k2 python bench.py --steps 1 --warmup 3 --profile-region --no-e2e --no-cpu-baseline --no-side --no-parity > gpurun_out/ncu_full.log 2>&1; echo "k2 rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off -c 3000 --csv --log-file gpurun_out/r02_e2e_launches_raw.csv python bench.py --steps 1 --warmup 3 --profile-e2e --batch 131072 --no-cpu-baseline --no-side --no-parity > gpurun_out/ncu_e2e.log 2>&1; echo "e2e rc=$?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:tsp_replay --launch-skip 1 -c 1 -f -o gpurun_out/r02_c2_tsp_replay_v2 python tools/profile_side.py c2 > gpurun_out/ncu_c2v2.log 2>&1; echo "c2 rc=$?"
ls -la gpurun_out | tail -12
