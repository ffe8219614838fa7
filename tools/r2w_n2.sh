#!/bin/bash
out=gpurun_out; mkdir -p $out
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29561 bench.py --gpus 2 --steps 5 > $out/r2w_bench_n2.json 2> $out/r2w_bench_n2.err; echo "n2 rc=$?"; tail -2 $out/r2w_bench_n2.err | cut -c1-200; cut -c1-200 $out/r2w_bench_n2.json
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29562 bench.py --gpus 2 --steps 5 --sharding samples > $out/r2w_bench_n2_samples.json 2> $out/r2w_bench_n2_samples.err; echo "n2 samples rc=$?"; cut -c1-200 $out/r2w_bench_n2_samples.json
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29563 bench.py --gpus 2 --impl reference --steps 1 --warmup 0 --cpu-rows 16 > $out/r2w_ref_n2.json 2> $out/r2w_ref_n2.err; echo "ref n2 rc=$?"; cut -c1-160 $out/r2w_ref_n2.json
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29564 tools/cfg5_sweep.py --out $out/r2w_sweep_check.json --spp 16 --grid-spp 2> $out/r2w_sweep.err; echo "sweep rc=$?"; tail -1 $out/r2w_sweep.err | cut -c1-160
timeout 300 python bench.py --inprocess 2 --steps 3 --no-cpu --lean > $out/r2w_inproc2.json 2> $out/r2w_inproc2.err; echo "inproc rc=$?"; cut -c1-160 $out/r2w_inproc2.json
