#!/bin/bash
# A/B of kernel tuning builds on one box: tools/ab.sh [bench args] -- lib1.so lib2.so ...
# Each library is benchmarked twice, interleaved, so box-to-box and clock drift cancel. Every run is under `timeout`:
# a kernel that does not terminate must not hold the box.
args=(); while [ $# -gt 0 ] && [ "$1" != "--" ]; do args+=("$1"); shift; done; shift
for rep in 1 2; do
  for lib in "$@"; do
    RTC_SKIP_ABI_CHECK=1 RTC_LIB_PATH=$PWD/$lib timeout 240 python bench.py --no-cpu --steps 2 "${args[@]}" 2>&1 | python -c "
import sys, json
txt = sys.stdin.read().strip().splitlines()
try:
    d = json.loads(txt[-1])
    print('%-50s %8.1f Mrays/s  %s' % ('$lib', d['value'], d.get('kernel_ms')))
except Exception:
    print('%-50s FAILED: %s' % ('$lib', ' | '.join(txt[-3:])[:300]))"
  done
done
