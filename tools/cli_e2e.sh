#!/bin/bash
# The whole program, file to file, as the reference's user runs it: bin/rtc on a generated atrium .glb (cfg3's scene,
# 1080p, 64 spp, 4 bounces), default path (stage1 assembly + stage2 on the device) and with stage1+2 on the host.
# tools/cli_e2e.sh <tag>
tag=${1:-rX}; out=$PWD/gpurun_out; mkdir -p $out
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
(cd $dir && $cli --in input.glb --out device.png --height 1080 --seed 1) 2> $out/${tag}_cli_device.log; echo "device rc=$?"
(cd $dir && RTC_HOST_BUILD=1 $cli --in input.glb --out host.png --height 1080 --seed 1) 2> $out/${tag}_cli_hostbuild.log; echo "host rc=$?"
cmp $dir/device.png $dir/host.png && echo "PNG files identical" | tee -a $out/${tag}_cli_device.log
rm -rf $dir
