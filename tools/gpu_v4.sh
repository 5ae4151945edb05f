#!/usr/bin/env bash
# v4 bring-up: parity of the v4 kernel, then bench v4 vs v3, then phase skipping (results wrong when phases are skipped)
OUT=gpurun_out; mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "v4 or baseline_sizes or full_size" > $OUT/v4_pytest.log 2>&1; tail -15 $OUT/v4_pytest.log
for v in 4 3; do
  timeout 300 python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu --fused-ver $v > $OUT/v4_bench_$v.json 2> $OUT/v4_bench_$v.err
  python -c "
import json; d=json.loads(open('$OUT/v4_bench_$v.json').read().strip().splitlines()[-1]); print('ver',$v,'ms',round(d['ms_per_step'],4),'U1/s',round(d['value']),'frac',round(d['roofline']['frac'],4))"
done
for m in 1 2 4 8 16 32 63; do
  CHB_FUSED_DBG=$m timeout 120 python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('dbg', $m, 'ms', round(d['ms_per_step'],4))"
done
