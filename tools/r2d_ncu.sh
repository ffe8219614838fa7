#!/bin/bash
out=gpurun_out; mkdir -p $out
python bench.py --steps 2 --no-cpu --accel bvh > $out/r2d_bench_bvh.json 2> $out/r2d_bench_bvh.err; echo "rc=$?"; cut -c1-200 $out/r2d_bench_bvh.json
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_trace_bvh|k_shade|k_gen|k_accum" -s 11 -c 10 -o $out/r2d_bvh \
    python bench.py --steps 1 --no-cpu --accel bvh > $out/r2d_ncu_full.log 2>&1; echo "ncu rc=$?"
