#!/usr/bin/env bash
# quick A/B of the bench kernel: parity of the v4 kernel, then ms per step (3 runs)
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "v4" 2>&1 | tail -1
for i in 1 2 3; do
timeout 120 python bench.py --steps 20 --warmup 3 --no-e2e --no-cpu --no-newton 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('ms', round(d['ms_per_step'],4), 'U1/s', round(d['value']), 'frac', round(d['roofline']['frac'],4))"
done
