#!/usr/bin/env bash
OUT=gpurun_out; mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "v4 or full_size" > $OUT/v4c_pytest.log 2>&1; tail -5 $OUT/v4c_pytest.log
run() { local label="$1"; shift
  env "$@" timeout 120 python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$label', 'ms', round(d['ms_per_step'],4), 'U1/s', round(d['value']))"
}
run "v4 stagger0" CHB_FUSED_STAGGER_NS=0
run "v4 stagger24k" CHB_FUSED_STAGGER_NS=24000
for m in 1 2 4 8 16 32 63; do run "dbg$m" CHB_FUSED_STAGGER_NS=0 CHB_FUSED_DBG=$m; done
