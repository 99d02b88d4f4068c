#!/bin/bash
# Where does the step time go?  Same bench with the optimizer overlapped / serial / skipped.
timeout 200 python tools/gpu_probe_corun.py 2>&1 | tail -7
for mode in "DLBB_DP_OVERLAP=1" "DLBB_DP_OVERLAP=0" "DLBB_DEBUG_SKIP_OPT=1" "DLBB_ADAM_OVERLAP_CTAS=148" "DLBB_ADAM_OVERLAP_CTAS=444"; do
  env $mode timeout 200 python bench.py --steps 8 --warmup 3 --no-e2e 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('$mode', round(d['ms_per_step'], 2), 'ms', d['clocks'])"
done
