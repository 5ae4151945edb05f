#!/bin/bash
# profiling aid: time of the fused kernel with phases skipped (CHB_FUSED_DBG bits: 1 load pass, 2 inverse passes, 4 pair,
# 8 forward passes, 16 store pass, 32 reduce, 64 = no tensor-memory stash); results are wrong when phases are skipped
for m in ${@:-0 1 2 4 8 16 32 63 64}; do
  CHB_FUSED_DBG=$m timeout 120 python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('dbg', $m, 'ms', round(d['ms_per_step'],4))"
done
