#!/bin/bash
# ncu --set full of the launches the bench TIMES (third frame: the whole frame is one batch from the second frame on), then the cfg5 sweep at N = 1
out=gpurun_out; mkdir -p $out
timeout 900 python tools/cfg5_sweep.py --out $out/r2u_cfg5_n1.json 2> $out/r2u_cfg5_n1.err; echo "sweep rc=$?"; tail -9 $out/r2u_cfg5_n1.err | cut -c1-150
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_trace_bvh|k_shade" -s 56 -c 8 -o $out/r2_final_bvh python bench.py --steps 1 --no-cpu --lean > $out/r2_final_ncu_bvh.log 2>&1; echo "ncu bvh rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_trace|k_shade" -s 56 -c 8 -o $out/r2_final_grid python bench.py --steps 1 --no-cpu --lean --accel grid > $out/r2_final_ncu_grid.log 2>&1; echo "ncu grid rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $out/r2_final_launches.csv python bench.py --steps 1 --no-cpu --lean > $out/r2_final_ncu_launch.log 2>&1; echo "ncu list rc=$?"
