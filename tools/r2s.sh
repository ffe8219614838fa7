#!/bin/bash
out=gpurun_out; mkdir -p $out
timeout 900 python -m pytest tests -m gpu -x -q > $out/r2s_tests.log 2>&1; echo "tests rc=$?"; tail -3 $out/r2s_tests.log
timeout 300 python bench.py --no-cpu --steps 5 > $out/r2s_bench.json 2> $out/r2s_bench.err; echo "bench rc=$?"; python -c "
import json; d=json.load(open('$out/r2s_bench.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['gpu_launches'], d['kernel_ms'], d.get('grid_exact'))"
timeout 300 python bench.py --no-cpu --steps 5 --accel grid > $out/r2s_bench_grid.json 2> $out/r2s_bench_grid.err; echo "grid rc=$?"; python -c "
import json; d=json.load(open('$out/r2s_bench_grid.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['gpu_launches'], d['kernel_ms'])"
timeout 300 python bench.py --workload cfg4_terrain5M_4k_256spp --spp 64 --steps 2 --no-cpu --lean > $out/r2s_cfg4_bvh.json 2> $out/r2s_cfg4_bvh.err; echo "cfg4 rc=$?"; python -c "
import json; d=json.load(open('$out/r2s_cfg4_bvh.json')); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['gpu_launches'], d['kernel_ms'])"
