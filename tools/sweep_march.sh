#!/bin/bash
# Rebuild the library with different occupancy targets for the march kernel and time C4.
for MINB in ${SWEEP:-3 4 5}; do
  make -s -B -C amrvolumerenderer_b200/csrc EXTRA_NVFLAGS="-DAMRVR_MARCH_MIN_BLOCKS=$MINB" > /dev/null 2>&1
  python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('MINB=$MINB', 'ms_per_step', round(d['ms_per_step'],4), 'march', round(d['config']['stage_ms']['march'],4), 'value', round(d['value'],1))
"
done
make -s -B -C amrvolumerenderer_b200/csrc > /dev/null 2>&1
