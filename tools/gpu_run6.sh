# round-2 GPU pass 6 (final, after the four-table tsp path and the enctc3 attention kernel): whole GPU suite, smoke, default bench,
# ncu evidence - launch list of the headline step, e2e list, full captures of the SLAP rollout kernel, the FP16 projection, the K2 TSP
# kernel (now forming the first-node context row itself) and the encoder attention kernel; cycle traces
set -x
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q -x --timeout 600 > gpurun_out/r02_pytest6.log 2>&1; echo "suite rc=$?"; tail -4 gpurun_out/r02_pytest6.log
timeout 300 python -m pytest tests/test_zz_no_alloc.py -q > gpurun_out/r02_no_alloc_alone.log 2>&1; echo "no-alloc alone rc=$?"
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02_smoke.log 2>&1; echo "smoke rc=$?"
timeout 1200 python bench.py > gpurun_out/r02_bench6.log 2> gpurun_out/r02_bench6.err; echo "bench rc=$?"; tail -c 300 gpurun_out/r02_bench6.log; tail -3 gpurun_out/r02_bench6.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off -c 600 --csv --log-file gpurun_out/r02_launches_raw.csv python bench.py --steps 2 --warmup 3 --profile-region --no-e2e --no-cpu-baseline --no-side --no-parity > gpurun_out/ncu_list.log 2>&1; echo "list rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off -c 3000 --csv --log-file gpurun_out/r02_e2e_launches_raw.csv python bench.py --steps 1 --warmup 3 --profile-e2e --batch 131072 --no-cpu-baseline --no-side --no-parity > gpurun_out/ncu_e2e.log 2>&1; echo "e2e rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:am_rollout_tsp_greedy --launch-skip 4 -c 1 -f -o gpurun_out/r02_k2 python bench.py --steps 1 --warmup 1 --batch 65536 --no-e2e --no-cpu-baseline --no-side --no-parity > gpurun_out/ncu_k2.log 2>&1; echo "k2 rc=$?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:am_rollout_slap --launch-skip 2 -c 1 -f -o gpurun_out/r02_c5_slap_v2 python tools/profile_side.py c5 > gpurun_out/ncu_c5.log 2>&1; echo "c5 rc=$?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:am_project_tc --launch-skip 2 -c 1 -f -o gpurun_out/r02_k3_f16 python tools/profile_side.py k3 4 > gpurun_out/ncu_k3.log 2>&1; echo "k3 rc=$?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:encoder_attention --launch-skip 2 -c 1 -f -o gpurun_out/r02_attention_tc3 python tools/profile_side.py attention > gpurun_out/ncu_att.log 2>&1; echo "attention rc=$?"
python tools/trace_slap.py build > /dev/null 2>&1; timeout 300 python tools/trace_slap.py > gpurun_out/r02_slap_trace.txt 2>&1; echo "trace rc=$?"
python tools/trace_rollout.py build > /dev/null 2>&1; timeout 300 python tools/trace_rollout.py > gpurun_out/r02_tsp_trace_5tables.txt 2>&1; RF_FC=1 timeout 300 python tools/trace_rollout.py > gpurun_out/r02_tsp_trace_4tables.txt 2>&1; echo "tsp trace rc=$?"
timeout 300 python tools/check_attention_tc.py > gpurun_out/r02_attention_check.txt 2>&1; echo "attention check rc=$?"
ls -la gpurun_out | tail -20
# same-box reference: the tree before the four-table path and enctc3 (commit 565e2ee unpacked and built under tools/_exp/old_tree by
#   git archive 565e2ee | tar -x -C tools/_exp/old_tree; (cd tools/_exp/old_tree; python -m rl4co_slap_b200.build) ), if present
if [ -d tools/_exp/old_tree ]; then (cd tools/_exp/old_tree && timeout 900 python bench.py --no-side --no-parity > /root/repo/gpurun_out/r02_bench6_prev_tree.log 2>/dev/null; echo "prev tree bench rc=$?"); fi
