#!/bin/bash
# last single-GPU visit of round 2: grid A/B of the queue prefetch, every -m gpu test, smoke, the bench lines with the final kernels
out=gpurun_out; mkdir -p $out
if [ -f build_variants/pf.so ]; then bash tools/ab.sh --lean --accel grid -- build_variants/pf.so build_variants/nopf.so > $out/r2z_ab_prefetch_grid.txt 2>&1; cut -c1-150 $out/r2z_ab_prefetch_grid.txt; fi
timeout 900 python -m pytest tests -m gpu -x -q > $out/r2z_final_tests.log 2>&1; echo "tests rc=$?"; tail -n 2 $out/r2z_final_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $out/r2z_final_smoke.log 2>&1; echo "smoke rc=$?"; tail -n 1 $out/r2z_final_smoke.log
timeout 600 python bench.py > $out/r2z_bench.json 2> $out/r2z_bench.err; echo "bench rc=$?"; cut -c1-200 $out/r2z_bench.json
timeout 600 python bench.py --accel grid > $out/r2z_bench_grid.json 2> $out/r2z_bench_grid.err; echo "grid rc=$?"; cut -c1-200 $out/r2z_bench_grid.json
timeout 600 python bench.py --workload cfg4_terrain5M_4k_256spp --spp 8 --steps 2 --no-cpu > $out/r2z_cfg4_bvh.json 2> $out/r2z_cfg4_bvh.err; echo "cfg4 rc=$?"; cut -c1-200 $out/r2z_cfg4_bvh.json
