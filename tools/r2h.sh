#!/bin/bash
out=gpurun_out; mkdir -p $out
bash tools/ab.sh --lean -- zig_raytracing_contest_b200/librtc.so build_variants/librtc_m5.so build_variants/librtc_c128.so 2>&1 | tee $out/r2h_ab_cfg3.txt
timeout 1500 python tools/cfg5_sweep.py --out $out/r2h_cfg5_n1.json 2> $out/r2h_cfg5_n1.err; echo "sweep rc=$?"; tail -12 $out/r2h_cfg5_n1.err
