# round-2 GPU pass 1: parity suite, bench with side configs, ncu captures of the side kernels
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x --timeout 600 > gpurun_out/r02_pytest1.log 2>&1; tail -15 gpurun_out/r02_pytest1.log
timeout 900 python bench.py > gpurun_out/r02_bench1.log 2> gpurun_out/r02_bench1.err; tail -c 1500 gpurun_out/r02_bench1.log; tail -5 gpurun_out/r02_bench1.err
timeout 300 ncu --set full --clock-control none --import-source on -k regex:tsp_replay --launch-skip 1 -c 1 -f -o gpurun_out/r02_c2_tsp_replay python tools/profile_side.py c2 > gpurun_out/ncu_c2.log 2>&1; tail -2 gpurun_out/ncu_c2.log
timeout 300 ncu --set full --clock-control none --import-source on -k regex:am_rollout_rf --launch-skip 1 -c 1 -f -o gpurun_out/r02_c5_slap python tools/profile_side.py c5 > gpurun_out/ncu_c5.log 2>&1; tail -2 gpurun_out/ncu_c5.log
timeout 300 ncu --set full --clock-control none --import-source on -k regex:am_rollout --launch-skip 1 -c 1 -f -o gpurun_out/r02_c4_pomo_simt python tools/profile_side.py c4 > gpurun_out/ncu_c4.log 2>&1; tail -2 gpurun_out/ncu_c4.log
ls -la gpurun_out | tail -8
