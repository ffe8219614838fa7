#!/bin/bash
# 8 GPUs, one process: rtc_render(num_gpus = 8) after the per-call device enumeration was removed; bin/rtc --gpus 8
out=$PWD/gpurun_out; mkdir -p $out
RTC_TRACE_MULTI=1 timeout 300 python bench.py --inprocess 8 --steps 3 --no-cpu --lean > $out/r2_n8_cfg3_inproc.json 2> $out/r2q_inproc8.err; echo "inproc rc=$?"; cut -c1-170 $out/r2_n8_cfg3_inproc.json; grep "rtc multi" $out/r2q_inproc8.err | tail -3 | tee $out/r2q_inproc8_host_phases.txt
timeout 300 python bench.py --inprocess 8 --steps 3 --no-cpu --lean --sharding samples > $out/r2_n8_cfg3_inproc_samples.json 2> $out/r2q_inproc8_samples.err; echo "inproc samples rc=$?"; cut -c1-170 $out/r2_n8_cfg3_inproc_samples.json
dir=$(mktemp -d)
python - "$dir" <<'PY'
import json, sys, os
sys.path.insert(0, os.getcwd())
from zig_raytracing_contest_b200 import gltf_io, scenes
d = sys.argv[1]
gltf_io.write_gltf(scenes.atrium(), os.path.join(d, "input.glb"), glb=True)
json.dump({"grid_resolution": [128, 128, 128], "num_threads": None, "num_samples": 64, "max_bounce": 4}, open(os.path.join(d, "config.json"), "w"))
PY
cli=$PWD/zig_raytracing_contest_b200/bin/rtc
(cd $dir && timeout 120 $cli --in input.glb --out one.png --height 1080 --seed 1) 2> $out/r2q_cli_gpus1.log; echo "cli 1 rc=$?"
(cd $dir && timeout 120 $cli --in input.glb --out eight.png --height 1080 --seed 1 --gpus 8) 2> $out/r2q_cli_gpus8.log; echo "cli 8 rc=$?"
(cd $dir && timeout 120 $cli --in input.glb --out eight_bvh.png --height 1080 --seed 1 --gpus 8 --accel bvh) 2> $out/r2q_cli_gpus8_bvh.log; echo "cli 8 bvh rc=$?"
cmp $dir/one.png $dir/eight.png && echo "PNG of --gpus 8 identical to --gpus 1" | tee -a $out/r2q_cli_gpus8.log
grep -E "Rendered|Rays|Done" $out/r2q_cli_gpus1.log $out/r2q_cli_gpus8.log $out/r2q_cli_gpus8_bvh.log
rm -rf $dir
