#!/bin/bash
out=gpurun_out; mkdir -p $out
timeout 600 python -m pytest tests/test_gpu_bvh.py -x -q > $out/r2c_tests_bvh.log 2>&1; echo "bvh tests rc=$?"; tail -3 $out/r2c_tests_bvh.log
bash tools/ab.sh --accel bvh -- zig_raytracing_contest_b200/librtc.so build_variants/librtc_bvh_m4.so build_variants/librtc_bvh_m4r4.so build_variants/librtc_bvh_m3r4.so build_variants/librtc_bvh_m4r8.so 2>&1 | tee $out/r2c_ab_cfg3.txt
bash tools/ab.sh --accel bvh --workload cfg4_terrain5M_4k_256spp --spp 8 -- zig_raytracing_contest_b200/librtc.so build_variants/librtc_bvh_m4.so 2>&1 | tee $out/r2c_ab_cfg4.txt
