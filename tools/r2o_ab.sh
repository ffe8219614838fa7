#!/bin/bash
out=gpurun_out; mkdir -p $out
timeout 300 python -m pytest tests/test_gpu_bvh.py -x -q -s > $out/r2o_tests.log 2>&1; echo "tests rc=$?"; grep -E "passed|failed|bvh vs oracle|bit-identical|Error|watchdog" $out/r2o_tests.log | tail -12
timeout 300 python -m pytest tests/test_gpu_fullsize.py -x -q -s > $out/r2o_tests_full.log 2>&1; echo "fullsize rc=$?"; grep -E "passed|failed|bvh vs oracle|bit-identical|Error|watchdog" $out/r2o_tests_full.log | tail -6
bash tools/ab.sh --lean -- zig_raytracing_contest_b200/librtc.so build_variants/librtc_p1.so build_variants/librtc_p8.so build_variants/librtc_p16.so build_variants/librtc_p24.so 2>&1 | tee $out/r2o_ab_cfg3.txt
bash tools/ab.sh --lean --workload cfg4_terrain5M_4k_256spp --spp 8 -- zig_raytracing_contest_b200/librtc.so build_variants/librtc_p1.so build_variants/librtc_p24.so 2>&1 | tee $out/r2o_ab_cfg4.txt
