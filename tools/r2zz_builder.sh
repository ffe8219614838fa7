#!/bin/bash
# the rewritten host BVH builder on the box: every -m gpu test through it, then old and new builder timed on the box's cores
out=gpurun_out; mkdir -p $out
timeout 60 python -m pytest tests -m gpu -x -q > $out/r2zz_tests.log 2>&1; echo "tests rc=$?"; tail -n 2 $out/r2zz_tests.log
RTC_TRACE_BVH=1 timeout 10 ./build_scratch/bvhtime 1581 16 > $out/r2zz_bvhtime_new.txt 2>&1; grep -v "depth 2" $out/r2zz_bvhtime_new.txt
timeout 10 ./build_scratch/bvhtime_old 1581 16 > $out/r2zz_bvhtime_old.txt 2>&1; cat $out/r2zz_bvhtime_old.txt
nproc
