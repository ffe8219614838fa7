#!/bin/bash
# One multi-GPU box visit (gpurun --gpus N): tools/gpu_multi.sh <tag> <N>
#   cfg3 default line (sample-sharded, weak), cfg4 (5M triangles, 4K, 256 spp, strips dealt over the ranks, strong),
#   cfg5 (spp sweep at 4K on the cfg4 scene, sample-sharded + NCCL reduce)
tag=${1:-rX}; N=${2:-8}
out=gpurun_out; mkdir -p $out
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N "$@"; }
run > $out/${tag}_n${N}_cfg3.json 2> $out/${tag}_n${N}_cfg3.err; echo "cfg3 rc=$?"
run --workload cfg4_terrain5M_4k_256spp --sharding tiles --steps 3 > $out/${tag}_n${N}_cfg4_tiles.json 2> $out/${tag}_n${N}_cfg4_tiles.err; echo "cfg4 rc=$?"
for total in ${SWEEP:-16 64 256 1024 4096}; do
  spp=$((total / N)); [ $spp -lt 1 ] && continue
  steps=2; [ $total -ge 1024 ] && steps=1
  run --workload cfg4_terrain5M_4k_256spp --spp $spp --steps $steps > $out/${tag}_n${N}_cfg5_spp${total}.json 2> $out/${tag}_n${N}_cfg5_spp${total}.err; echo "cfg5 $total rc=$?"
done
echo done
