#!/usr/bin/env bash
# round 2, run U: tiled pair kernel with balanced rounds; transform split / block-size A/B with the batched loads
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_api2.py tests/test_big_split.py tests/test_ode_forces.py -m gpu -x -q > $OUT/r2u_pytest.txt 2>&1; echo "pytest exit $?" >> $OUT/r2u_pytest.txt
tail -3 $OUT/r2u_pytest.txt
run() { local label="$1"; shift
  timeout 300 python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu --no-newton --no-configs --stages "$@" 2>$OUT/cfg_err.txt | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); s=d.get('stages') or {}
print('$label', 'ms', round(d['ms_per_step'],4), 'U1/s', round(d['value']), 'fusedver', d.get('fused_version'), 'launches', d['gpu_launches'], {k: round(v,4) for k,v in s.items()})" || tail -3 $OUT/cfg_err.txt
}
{
run "C4 64C@16384 b256" --system 64C --nint 16384 --batch 256
run "C4 64C@16384 b1024" --system 64C --nint 16384 --batch 1024
for knobs in "CHB_BIG_R=4 CHB_BIG_T=256" "CHB_BIG_R=8 CHB_BIG_T=256" "CHB_BIG_R=4" "CHB_BIG_R=2 CHB_BIG_T=256"; do
  env $knobs bash -c "$(declare -f run); OUT=$OUT; run 'C4 64C@16384 b256 [$knobs]' --system 64C --nint 16384 --batch 256"
done
run "16C@4096 b1024" --system 16C --nint 4096 --batch 1024
run "20B@1200 b1024" --system 20B --nint 1200 --batch 1024
} > $OUT/r2u_configs.txt 2>&1
cat $OUT/r2u_configs.txt
CMD="python bench.py --system 64C --nint 16384 --batch 256 --steps 2 --warmup 3 --no-e2e --no-cpu --no-newton --no-configs"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file $OUT/r2u_launches_64C_b256.csv $CMD > $OUT/r2u_ncu.log 2>&1
grep -v "^==" $OUT/r2u_launches_64C_b256.csv | tail -5 | cut -c1-200
