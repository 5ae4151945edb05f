#!/usr/bin/env bash
OUT=gpurun_out; mkdir -p $OUT
run() { # label, env...
  local label="$1"; shift
  env "$@" timeout 120 python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$label', 'ms', round(d['ms_per_step'],4), 'U1/s', round(d['value']))"
}
for s in 0 8000 16000 24000 32000 48000; do run "stagger=$s" CHB_FUSED_STAGGER_NS=$s; done
run "stagger0+nofence" CHB_FUSED_STAGGER_NS=0 CHB_FUSED_DBG=1024
run "stagger0+notmemld" CHB_FUSED_STAGGER_NS=0 CHB_FUSED_DBG=2048
run "stagger0+noprefetch" CHB_FUSED_STAGGER_NS=0 CHB_FUSED_DBG=512
run "stagger24k+noprefetch" CHB_FUSED_STAGGER_NS=24000 CHB_FUSED_DBG=512
run "stagger24k+nostorepass" CHB_FUSED_STAGGER_NS=24000 CHB_FUSED_DBG=16
run "stagger24k+allskipped" CHB_FUSED_STAGGER_NS=24000 CHB_FUSED_DBG=63
