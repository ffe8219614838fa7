#!/bin/bash
# round 2, 8 GPUs: the driver's SCALE shape (cfg3, strong) through rtc_render_sharded, the one-process path, cfg4 at its full 256 spp
out=gpurun_out; mkdir -p $out
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $1 bench.py --gpus 8 "${@:3}" > $out/$2.json 2> $out/$2.err; echo "$2 rc=$?"; cut -c1-160 $out/$2.json; }
run 29531 r2_n8_cfg3_strips --steps 5
timeout 600 python bench.py --inprocess 8 --steps 3 --no-cpu --lean > $out/r2_n8_cfg3_inproc.json 2> $out/r2_n8_cfg3_inproc.err; echo "inproc rc=$?"; cut -c1-160 $out/r2_n8_cfg3_inproc.json
run 29533 r2_n8_cfg4_strips --steps 2 --workload cfg4_terrain5M_4k_256spp --lean
timeout 300 python -m pytest tests/test_gpu_multi.py -x -q > $out/r2_n8_tests_multi.log 2>&1; echo "multi tests rc=$?"; tail -2 $out/r2_n8_tests_multi.log
