#!/bin/bash
# Run bench.py (cfg1, short) for each variant library under _exp/ named on the command line; prints one line each.
cd "$(dirname "$0")/.."
for v in "$@"; do
  if [ "$v" = base ]; then unset MTSB200_LIB; else export MTSB200_LIB=$PWD/_exp/$v.so; fi
  python bench.py --steps 2 --warmup 2 --no-cpu-baseline ${BENCH_ARGS} 2>/dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.readlines()[-1]); s = d['stage_ms_last_step']
print('$v', 'ms/step %.1f' % d['ms_per_step'], ' '.join('%s=%.1f' % (k, v) for k, v in s.items()))"
done
