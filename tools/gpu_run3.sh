# round-2 GPU pass 3: whole GPU suite, bench with side configs (C4 now on the tensor-core multistart kernel), ncu of that kernel
set -x
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q -x --timeout 600 > gpurun_out/r02_pytest3.log 2>&1; echo "suite rc=$?"; tail -5 gpurun_out/r02_pytest3.log
timeout 900 python bench.py > gpurun_out/r02_bench2.log 2> gpurun_out/r02_bench2.err; echo "bench rc=$?"; tail -c 600 gpurun_out/r02_bench2.log; tail -5 gpurun_out/r02_bench2.err
timeout 300 ncu --set full --clock-control none --import-source on -k regex:am_rollout_ms_tc --launch-skip 1 -c 1 -f -o gpurun_out/r02_c4_pomo_tc python tools/profile_side.py c4 > gpurun_out/ncu_c4tc.log 2>&1; tail -2 gpurun_out/ncu_c4tc.log
ls -la gpurun_out | tail -6
