#!/bin/bash
# ncu --set full of k_shade, bounces 0-3 of a ONE-BATCH frame (launches 28..31 of k_shade: frame 1 = 7 batches x 4 bounces), with the L2 prefetch
out=gpurun_out; mkdir -p $out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_shade -s 28 -c 4 -o /tmp/r2z_shade python bench.py --steps 1 --no-cpu --lean > $out/r2z_ncu_shade.log 2>&1; echo "ncu shade rc=$?"
python tools/ncu_summary.py /tmp/r2z_shade.ncu-rep "r2z k_shade<0,1> (BVH path), bounces 0-3 of a ONE-BATCH frame of cfg3, RTC_SHADE_PREFETCH=2 (bench.py --steps 1 --no-cpu --lean, launches 28-31)" > $out/r2z_ncu_full_shade_onebatch.txt 2>&1
cut -c1-160 $out/r2z_ncu_full_shade_onebatch.txt
sz=$(stat -c %s /tmp/r2z_shade.ncu-rep 2>/dev/null || echo 0); if [ "$sz" -gt 0 ] && [ "$sz" -lt 30000000 ]; then cp /tmp/r2z_shade.ncu-rep $out/; fi
