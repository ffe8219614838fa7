#!/bin/bash
out=$PWD/gpurun_out; mkdir -p $out
timeout 300 python -m pytest tests/test_gpu_multi.py -x -q > $out/r2r_tests_multi.log 2>&1; echo "multi tests rc=$?"; tail -2 $out/r2r_tests_multi.log
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
(cd $dir && timeout 120 $cli --in input.glb --out one.png --height 1080 --seed 1) 2> $out/r2r_cli_gpus1.log; echo "cli 1 rc=$?"
(cd $dir && timeout 120 $cli --in input.glb --out two.png --height 1080 --seed 1 --gpus 2) 2> $out/r2r_cli_gpus2.log; echo "cli 2 rc=$?"
(cd $dir && timeout 120 $cli --in input.glb --out two_bvh.png --height 1080 --seed 1 --gpus 2 --accel bvh) 2> $out/r2r_cli_gpus2_bvh.log; echo "cli 2 bvh rc=$?"
cmp $dir/one.png $dir/two.png && echo "PNG of --gpus 2 identical to --gpus 1" | tee -a $out/r2r_cli_gpus2.log
grep -E "Prepared|Rendered|Rays|Done" $out/r2r_cli_gpus1.log $out/r2r_cli_gpus2.log $out/r2r_cli_gpus2_bvh.log
rm -rf $dir
