#!/bin/bash
out=gpurun_out; mkdir -p $out
python bench.py --steps 2 --no-cpu --lean > $out/r2k_bench.json 2> $out/r2k_bench.err; echo "rc=$?"; cut -c1-120 $out/r2k_bench.json
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_trace_bvh|k_shade" -s 10 -c 8 -o $out/r2k_bvh \
    python bench.py --steps 1 --no-cpu --lean > $out/r2k_ncu_full.log 2>&1; echo "ncu rc=$?"
