#!/usr/bin/env bash
OUT=gpurun_out; mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "v4 or full_size" > $OUT/v4d_pytest.log 2>&1; tail -3 $OUT/v4d_pytest.log
run() { local label="$1"; shift
  env "$@" timeout 120 python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$label', 'ms', round(d['ms_per_step'],4), 'U1/s', round(d['value']))"
}
run "v4" A=1
run "v4 noprefetch" CHB_FUSED_DBG=512
run "v4 stagger16k" CHB_FUSED_STAGGER_NS=16000
for b in 1024 8192 16384; do
timeout 120 python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu --batch $b 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('batch $b', 'ms', round(d['ms_per_step'],4), 'U1/s', round(d['value']))"
done
