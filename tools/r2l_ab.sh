#!/bin/bash
out=gpurun_out; mkdir -p $out
timeout 600 python -m pytest tests/test_gpu_bvh.py -x -q > $out/r2l_tests.log 2>&1; echo "tests rc=$?"; tail -2 $out/r2l_tests.log
bash tools/ab.sh --lean -- zig_raytracing_contest_b200/librtc.so build_variants/librtc_top0.so build_variants/librtc_top9.so build_variants/librtc_top256.so 2>&1 | tee $out/r2l_ab_cfg3.txt
bash tools/ab.sh --lean --workload cfg4_terrain5M_4k_256spp --spp 8 -- zig_raytracing_contest_b200/librtc.so build_variants/librtc_top0.so build_variants/librtc_top256.so 2>&1 | tee $out/r2l_ab_cfg4.txt
