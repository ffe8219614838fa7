#!/bin/bash
out=gpurun_out; mkdir -p $out
timeout 900 python -m pytest tests/test_gpu_bvh.py tests/test_gpu_parity.py tests/test_gpu_build.py tests/test_gpu_gltf.py -x -q > $out/r2g_tests.log 2>&1; echo "tests rc=$?"; tail -3 $out/r2g_tests.log
bash tools/ab.sh --lean -- zig_raytracing_contest_b200/librtc.so build_variants/librtc_d0.so build_variants/librtc_d2.so build_variants/librtc_d6.so build_variants/librtc_s4.so 2>&1 | tee $out/r2g_ab_cfg3.txt
bash tools/ab.sh --lean --workload cfg4_terrain5M_4k_256spp --spp 8 -- zig_raytracing_contest_b200/librtc.so build_variants/librtc_d0.so build_variants/librtc_s4.so 2>&1 | tee $out/r2g_ab_cfg4.txt
