#!/bin/bash
out=gpurun_out; mkdir -p $out
timeout 900 python -m pytest tests -m gpu -x -q -s > $out/r2_final_tests.log 2>&1; echo "tests rc=$?"; grep -E "passed|failed|cfg3 full frame|Error" $out/r2_final_tests.log | tail -6
