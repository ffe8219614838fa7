#!/bin/bash
out=gpurun_out; mkdir -p $out
timeout 900 python -m pytest tests/test_gpu_bvh.py tests/test_gpu_parity.py -x -q -s > $out/r2j_tests.log 2>&1; echo "tests rc=$?"; grep -E "fast math|passed|failed" $out/r2j_tests.log | tail -5
bash tools/ab.sh --lean -- zig_raytracing_contest_b200/librtc.so build_variants/librtc_nosort.so build_variants/librtc_nold256.so 2>&1 | tee $out/r2j_ab_cfg3.txt
bash tools/ab.sh --lean --accel grid -- zig_raytracing_contest_b200/librtc.so build_variants/librtc_nosort.so 2>&1 | tee $out/r2j_ab_cfg3_grid.txt
bash tools/ab.sh --lean --workload cfg4_terrain5M_4k_256spp --spp 8 -- zig_raytracing_contest_b200/librtc.so build_variants/librtc_nosort.so build_variants/librtc_nold256.so 2>&1 | tee $out/r2j_ab_cfg4.txt
