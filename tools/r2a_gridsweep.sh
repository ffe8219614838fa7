#!/bin/bash
# round 2, visit a: baseline line, --math fast line, and the exact grid path on cfg4 at several grid_resolution values
out=gpurun_out; mkdir -p $out
python bench.py --steps 3 > $out/r2a_bench.json 2> $out/r2a_bench.err; echo "bench rc=$?"
python bench.py --steps 3 --math fast --no-cpu > $out/r2a_bench_fast.json 2> $out/r2a_bench_fast.err; echo "fast rc=$?"
for g in "128 128 128" "256 64 256" "512 64 512" "512 128 512" "1024 64 1024" "1024 128 1024"; do
  n=$(echo $g | tr ' ' 'x')
  timeout 600 python bench.py --workload cfg4_terrain5M_4k_256spp --spp 8 --steps 1 --no-cpu --grid $g > $out/r2a_cfg4_$n.json 2> $out/r2a_cfg4_$n.err; echo "cfg4 $n rc=$?"
done
for g in "192 64 96" "256 96 128" "256 128 256"; do
  n=$(echo $g | tr ' ' 'x')
  timeout 300 python bench.py --spp 16 --steps 2 --no-cpu --grid $g > $out/r2a_cfg3_$n.json 2> $out/r2a_cfg3_$n.err; echo "cfg3 $n rc=$?"
done
echo done
