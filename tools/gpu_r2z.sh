#!/usr/bin/env bash
# round 2, run Z: register / sub-block variants of the tiled pair kernel (A/B), 64C and 16C
OUT=gpurun_out; mkdir -p $OUT
run() { local label="$1"; shift
  timeout 300 python bench.py --steps 20 --warmup 3 --no-e2e --no-cpu --no-newton --no-configs "$@" 2>$OUT/cfg_err.txt | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$label', 'ms', round(d['ms_per_step'],4), 'U1/s', round(d['value']))" || tail -3 $OUT/cfg_err.txt
}
{
for v in 0 1 2 3; do
  env CHB_TILE_VAR=$v bash -c "$(declare -f run); OUT=$OUT; run 'C4 64C@16384 b256 [tile variant $v]' --system 64C --nint 16384 --batch 256"
  env CHB_TILE_VAR=$v bash -c "$(declare -f run); OUT=$OUT; run '16C@4096 b1024 [tile variant $v]' --system 16C --nint 4096 --batch 1024"
done
} > $OUT/r2z_tile_variants.txt 2>&1
cat $OUT/r2z_tile_variants.txt
