#!/usr/bin/env bash
# round 2, run V: start offset of the first-wave CTAs on odd SMs in the staged transforms (A/B), 64C
OUT=gpurun_out; mkdir -p $OUT
run() { local label="$1"; shift
  timeout 300 python bench.py --steps 20 --warmup 3 --no-e2e --no-cpu --no-newton --no-configs "$@" 2>$OUT/cfg_err.txt | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$label', 'ms', round(d['ms_per_step'],4), 'U1/s', round(d['value']))" || tail -3 $OUT/cfg_err.txt
}
{
for ns in 0 4000 8000 12000 16000 24000; do
  env CHB_BIG_STAGGER_NS=$ns bash -c "$(declare -f run); OUT=$OUT; run 'C4 64C@16384 b256 [stagger $ns ns]' --system 64C --nint 16384 --batch 256"
done
for ns in 0 8000 16000; do
  env CHB_BIG_STAGGER_NS=$ns bash -c "$(declare -f run); OUT=$OUT; run 'C4 64C@16384 b1024 [stagger $ns ns]' --system 64C --nint 16384 --batch 1024"
  env CHB_BIG_STAGGER_NS=$ns bash -c "$(declare -f run); OUT=$OUT; run '16C@4096 b1024 [stagger $ns ns]' --system 16C --nint 4096 --batch 1024"
done
} > $OUT/r2v_stagger.txt 2>&1
cat $OUT/r2v_stagger.txt
