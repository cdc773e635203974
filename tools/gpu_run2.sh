# round-2 GPU pass 2: the no-alloc entry point alone in a fresh process, then the whole GPU suite (it collects last)
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x --timeout 300 > gpurun_out/r02_zz_alone.log 2>&1; echo "alone rc=$?"; tail -15 gpurun_out/r02_zz_alone.log
timeout 1800 python -m pytest tests -m gpu -q -x --timeout 600 > gpurun_out/r02_pytest2.log 2>&1; echo "suite rc=$?"; tail -15 gpurun_out/r02_pytest2.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r02_smoke.log 2>&1; echo "smoke rc=$?"; tail -4 gpurun_out/r02_smoke.log
