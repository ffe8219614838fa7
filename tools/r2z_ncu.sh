#!/bin/bash
# ncu --set full of the traversal launches the bench TIMES: from a scene's second frame on the whole frame is one batch
# (two-stage scratch budget), so k_trace_bvh launches 28..31 (frame 1 = 7 batches x 4 bounces) are bounces 0-3 of frame 2.
# The report stays on the box (gpurun_out/ is capped at 64 MiB); the table of tools/ncu_summary.py comes back.
out=gpurun_out; mkdir -p $out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_trace_bvh -s 28 -c 4 -o /tmp/r2z_bvh python bench.py --steps 1 --no-cpu --lean > $out/r2z_ncu_bvh.log 2>&1; echo "ncu bvh rc=$?"
python tools/ncu_summary.py /tmp/r2z_bvh.ncu-rep "r2z BVH path: k_trace_bvh, bounces 0-3 of a ONE-BATCH frame of cfg3 (132.7 M paths; bench.py --steps 1 --no-cpu --lean, launches 28-31)" > $out/r2z_ncu_full_bvh_onebatch.txt 2>&1
cat $out/r2z_ncu_full_bvh_onebatch.txt | cut -c1-160
ls -la /tmp/r2z_bvh.ncu-rep
sz=$(stat -c %s /tmp/r2z_bvh.ncu-rep 2>/dev/null || echo 0); if [ "$sz" -gt 0 ] && [ "$sz" -lt 30000000 ]; then cp /tmp/r2z_bvh.ncu-rep $out/; fi
