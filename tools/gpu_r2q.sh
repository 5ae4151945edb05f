#!/usr/bin/env bash
# round 2, run Q: staged transforms with batched loads + slot table in shared memory: parity tests, A/B on the staged configurations, deep-level Newton probe, launch times
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests/test_big_split.py tests/test_gpu_parity.py tests/test_newton.py tests/test_zz_no_alloc.py -m gpu -x -q > $OUT/r2q_pytest.txt 2>&1; echo "pytest exit $?" >> $OUT/r2q_pytest.txt
tail -4 $OUT/r2q_pytest.txt
run() { local label="$1"; shift
  timeout 300 python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu --no-newton --no-configs --stages "$@" 2>$OUT/cfg_err.txt | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); s=d.get('stages') or {}
print('$label', 'ms', round(d['ms_per_step'],4), 'U1/s', round(d['value']), 'frac', round(d['roofline']['frac'],4), 'fusedver', d.get('fused_version'), 'launches', d['gpu_launches'], {k: round(v,4) for k,v in s.items()})" || tail -3 $OUT/cfg_err.txt
}
{
for knobs in "CHB_BIG_R=2" "CHB_BIG_R=4" "CHB_BIG_R=8"; do
  env $knobs bash -c "$(declare -f run); OUT=$OUT; run 'C4 64C@16384 b256 [$knobs]' --system 64C --nint 16384 --batch 256"
done
run "C4 64C@16384 b1024" --system 64C --nint 16384 --batch 1024
run "16C@4096 b1024" --system 16C --nint 4096 --batch 1024
run "20B@1200 b1024" --system 20B --nint 1200 --batch 1024
run "3C@49152 b16" --system 3C --nint 49152 --batch 16
run "3C@24576 b32" --system 3C --nint 24576 --batch 32
} > $OUT/r2q_configs.txt 2>&1
cat $OUT/r2q_configs.txt
timeout 300 python tools/newton_deep_probe.py --reps 2 > $OUT/r2q_probe.txt 2>&1; cat $OUT/r2q_probe.txt
timeout 600 python tools/sweep_bench.py --guesses 8192 --chunk 1024 --workers 4 > $OUT/r2q_sweep_8192_w4.json 2> $OUT/r2q_sweep_err.txt; head -c 500 $OUT/r2q_sweep_8192_w4.json; tail -2 $OUT/r2q_sweep_err.txt
