#!/bin/bash
# the allocation-free entry point: its test alone in a fresh process, then the whole -m gpu suite (it is collected last), then smoke()
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/r2z_noalloc_alone.log 2>&1; echo "alone rc=$?" | tee -a gpurun_out/r2z_noalloc_alone.log
timeout 600 python -m pytest tests -x -q -m gpu > gpurun_out/r2z_tests.log 2>&1; echo "suite rc=$?" | tee -a gpurun_out/r2z_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r2z_smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/r2z_smoke.log
tail -n 5 gpurun_out/r2z_noalloc_alone.log gpurun_out/r2z_tests.log gpurun_out/r2z_smoke.log
