#!/bin/bash
# round 2, visit e (2 GPUs): the multi-GPU C ABI — tests, then bench at N = 1 and N = 2 (torchrun) and in-process
out=gpurun_out; mkdir -p $out
nvidia-smi -L | head -3
timeout 900 python -m pytest tests/test_gpu_multi.py -x -q > $out/r2e_tests_multi.log 2>&1; echo "multi tests rc=$?"; tail -25 $out/r2e_tests_multi.log
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_bvh.py -x -q > $out/r2e_tests_parity.log 2>&1; echo "parity tests rc=$?"; tail -5 $out/r2e_tests_parity.log
timeout 600 python bench.py --steps 3 --no-cpu > $out/r2e_bench_n1.json 2> $out/r2e_bench_n1.err; echo "n1 rc=$?"; tail -3 $out/r2e_bench_n1.err; cut -c1-300 $out/r2e_bench_n1.json
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 3 > $out/r2e_bench_n2.json 2> $out/r2e_bench_n2.err; echo "n2 rc=$?"; tail -5 $out/r2e_bench_n2.err; cut -c1-300 $out/r2e_bench_n2.json
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 3 --sharding samples > $out/r2e_bench_n2_samples.json 2> $out/r2e_bench_n2_samples.err; echo "n2 samples rc=$?"; tail -5 $out/r2e_bench_n2_samples.err; cut -c1-300 $out/r2e_bench_n2_samples.json
timeout 600 python bench.py --steps 3 --no-cpu --inprocess 2 > $out/r2e_bench_inproc2.json 2> $out/r2e_bench_inproc2.err; echo "inproc rc=$?"; tail -3 $out/r2e_bench_inproc2.err; cut -c1-300 $out/r2e_bench_inproc2.json
echo done
