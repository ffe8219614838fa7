N=8; out=gpurun_out; tag=r1y
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N "$@"; }
for total in 16 64 1024; do
  spp=$((total / N)); steps=2; [ $total -ge 1024 ] && steps=1
  run --workload cfg4_terrain5M_4k_256spp --spp $spp --steps $steps > $out/${tag}_n${N}_cfg5_spp${total}.json 2> $out/${tag}_n${N}_cfg5_spp${total}.err; echo "cfg5 $total rc=$?"
done
