#!/bin/bash
# One GPU-box visit: parity tests, the default bench line, the cfg4 scene (HBM-sized) at a reduced spp,
# then the ncu launch list and one --set full capture of k_trace ON THE DEFAULT BENCH WORKLOAD (so the
# captured launches are the launches bench.py times). tools/gpu_round.sh <tag>
tag=${1:-rX}
out=gpurun_out
mkdir -p $out
python -m pytest tests -m gpu -x -q > $out/${tag}_tests.log 2>&1; echo "tests rc=$?" >> $out/${tag}_tests.log
tail -3 $out/${tag}_tests.log
python bench.py > $out/${tag}_bench.json 2> $out/${tag}_bench.err; echo "bench rc=$?"
python bench.py --workload cfg4_terrain5M_4k_256spp --spp 4 --steps 2 --no-cpu > $out/${tag}_bench_cfg4.json 2> $out/${tag}_bench_cfg4.err; echo "cfg4 rc=$?"
python bench.py --steps 1 --no-cpu > $out/${tag}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 500 --csv --log-file $out/${tag}_launches.csv \
    python bench.py --steps 1 --no-cpu > $out/${tag}_ncu_launch.log 2>&1
# second frame of the run (warm-up step 2), first batch: its four k_trace launches (bounce 0..3)
ncu --set full --clock-control none --import-source on -k regex:k_trace -s 24 -c 4 -o $out/${tag}_ktrace \
    python bench.py --steps 1 --no-cpu > $out/${tag}_ncu_full.log 2>&1
python tools/stage1_bench.py --tris 2000000 --out $out/${tag}_stage1.json > $out/${tag}_stage1.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"k_assemble|k_texels" -c 6 -o $out/${tag}_kassemble \
    python tools/stage1_bench.py --tris 2000000 --reps 1 > $out/${tag}_ncu_kassemble.log 2>&1
python tools/png_bench.py --out $out/${tag}_png.json > /dev/null 2>&1
echo done
