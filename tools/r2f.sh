#!/bin/bash
out=gpurun_out; mkdir -p $out
timeout 1200 python -m pytest tests/test_gpu_bvh.py tests/test_gpu_parity.py tests/test_gpu_fullsize.py -x -q -s > $out/r2f_tests.log 2>&1; echo "tests rc=$?"; grep -E "cfg4 5M|passed|failed|Error" $out/r2f_tests.log | tail -12
timeout 600 python bench.py --steps 3 --no-cpu > $out/r2f_bench.json 2> $out/r2f_bench.err; echo "rc=$?"; tail -2 $out/r2f_bench.err; cut -c1-250 $out/r2f_bench.json
timeout 600 python bench.py --steps 3 --no-cpu --accel grid > $out/r2f_bench_grid.json 2> $out/r2f_bench_grid.err; echo "rc=$?"; cut -c1-250 $out/r2f_bench_grid.json
timeout 600 python bench.py --workload cfg4_terrain5M_4k_256spp --spp 8 --steps 1 --no-cpu --lean > $out/r2f_cfg4_bvh.json 2> $out/r2f_cfg4_bvh.err; echo "cfg4 rc=$?"; cut -c1-250 $out/r2f_cfg4_bvh.json
ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file $out/r2f_launches.csv python bench.py --steps 1 --no-cpu --lean > $out/r2f_ncu_launch.log 2>&1; echo "ncu list rc=$?"
