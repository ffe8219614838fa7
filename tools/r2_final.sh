#!/bin/bash
# round 2, final single-GPU visit: every -m gpu test, smoke, the default bench line + the grid line + the reference arm,
# the ncu launch list of the bench command and one --set full capture of the traversal + shading kernels of each path
out=gpurun_out; mkdir -p $out
timeout 1800 python -m pytest tests -m gpu -x -q -s > $out/r2_final_tests.log 2>&1; echo "tests rc=$?"; grep -E "passed|failed|skipped|bvh vs oracle|bit-identical|fast math" $out/r2_final_tests.log | tail -20
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $out/r2_final_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 $out/r2_final_smoke.log
timeout 900 python bench.py > $out/r2_final_bench.json 2> $out/r2_final_bench.err; echo "bench rc=$?"; cut -c1-160 $out/r2_final_bench.json
timeout 900 python bench.py --accel grid > $out/r2_final_bench_grid.json 2> $out/r2_final_bench_grid.err; echo "grid rc=$?"; cut -c1-160 $out/r2_final_bench_grid.json
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > $out/r2_final_reference.json 2> $out/r2_final_reference.err; echo "ref rc=$?"; cut -c1-160 $out/r2_final_reference.json
timeout 600 python bench.py --workload cfg4_terrain5M_4k_256spp --spp 8 --steps 2 --no-cpu > $out/r2_final_cfg4_bvh.json 2> $out/r2_final_cfg4_bvh.err; echo "cfg4 rc=$?"; cut -c1-160 $out/r2_final_cfg4_bvh.json
timeout 600 python bench.py --steps 1 --no-cpu --lean > $out/r2_final_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $out/r2_final_launches.csv python bench.py --steps 1 --no-cpu --lean > $out/r2_final_ncu_launch.log 2>&1; echo "ncu list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_trace_bvh|k_shade" -s 10 -c 8 -o $out/r2_final_bvh python bench.py --steps 1 --no-cpu --lean > $out/r2_final_ncu_bvh.log 2>&1; echo "ncu bvh rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_trace|k_shade" -s 10 -c 8 -o $out/r2_final_grid python bench.py --steps 1 --no-cpu --lean --accel grid > $out/r2_final_ncu_grid.log 2>&1; echo "ncu grid rc=$?"
echo done
