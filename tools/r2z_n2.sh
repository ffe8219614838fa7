#!/bin/bash
# 2 GPUs with the final kernels: the multi-GPU tests (skipped on one-GPU boxes) and the strong-scaling bench line through rtc_render_sharded
out=gpurun_out; mkdir -p $out
timeout 200 python -m pytest tests/test_gpu_multi.py tests/test_zz_no_alloc.py -x -q -m gpu > $out/r2z_n2_tests_multi.log 2>&1; echo "multi tests rc=$?"; tail -n 2 $out/r2z_n2_tests_multi.log
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29571 bench.py --gpus 2 --steps 5 > $out/r2z_bench_n2.json 2> $out/r2z_bench_n2.err; echo "n2 rc=$?"; cut -c1-200 $out/r2z_bench_n2.json
