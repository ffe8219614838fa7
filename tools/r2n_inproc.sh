#!/bin/bash
out=gpurun_out; mkdir -p $out
N=${1:-4}
RTC_TRACE_MULTI=1 timeout 600 python bench.py --inprocess $N --steps 3 --no-cpu --lean > $out/r2n_inproc$N.json 2> $out/r2n_inproc$N.err; echo "inproc rc=$?"; cut -c1-200 $out/r2n_inproc$N.json; grep "rtc multi" $out/r2n_inproc$N.err | tail -9
timeout 600 python bench.py --inprocess $N --steps 3 --no-cpu --lean --sharding samples > $out/r2n_inproc${N}_samples.json 2> $out/r2n_inproc${N}_samples.err; echo "inproc samples rc=$?"; cut -c1-200 $out/r2n_inproc${N}_samples.json
timeout 300 python -m pytest tests/test_gpu_multi.py -x -q > $out/r2n_tests_multi.log 2>&1; echo "multi tests rc=$?"; tail -2 $out/r2n_tests_multi.log
