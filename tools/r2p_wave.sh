#!/bin/bash
out=gpurun_out; mkdir -p $out
for b in 6442450944 12884901888 25769803776 51539607552; do
  RTC_WAVE_BYTES=$b timeout 240 python bench.py --no-cpu --lean --steps 3 2> $out/r2p_wave_$b.err | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1]); print('RTC_WAVE_BYTES=$b', d['value'], 'Mrays/s', d['ms_per_step'], 'ms', 'launches', d['gpu_launches'], d['kernel_ms'])" | tee -a $out/r2p_wave.txt
done
for b in 6442450944 25769803776; do
  RTC_WAVE_BYTES=$b timeout 240 python bench.py --no-cpu --lean --steps 3 --accel grid 2>> $out/r2p_wave_$b.err | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1]); print('grid RTC_WAVE_BYTES=$b', d['value'], 'Mrays/s', d['ms_per_step'], 'ms', 'launches', d['gpu_launches'], d['kernel_ms'])" | tee -a $out/r2p_wave.txt
done
