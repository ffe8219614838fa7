#!/bin/bash
# round 2, visit b: first GPU run of RTC_ACCEL_BVH — parity tests, bench lines, launch list + one full ncu capture
out=gpurun_out; mkdir -p $out
timeout 900 python -m pytest tests/test_gpu_bvh.py -x -q -s > $out/r2b_tests_bvh.log 2>&1; echo "bvh tests rc=$?"; tail -15 $out/r2b_tests_bvh.log
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q > $out/r2b_tests_parity.log 2>&1; echo "parity tests rc=$?"; tail -3 $out/r2b_tests_parity.log
timeout 600 python bench.py --steps 3 --accel bvh --no-cpu > $out/r2b_bench_bvh.json 2> $out/r2b_bench_bvh.err; echo "bvh bench rc=$?"; cat $out/r2b_bench_bvh.json | cut -c1-600
timeout 600 python bench.py --workload cfg4_terrain5M_4k_256spp --spp 8 --steps 1 --no-cpu --accel bvh > $out/r2b_cfg4_bvh.json 2> $out/r2b_cfg4_bvh.err; echo "cfg4 bvh rc=$?"; cat $out/r2b_cfg4_bvh.json | cut -c1-600
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_trace_bvh -s 8 -c 4 -o $out/r2b_ktrace_bvh \
    python bench.py --steps 1 --no-cpu --accel bvh > $out/r2b_ncu_full.log 2>&1; echo "ncu rc=$?"
echo done
