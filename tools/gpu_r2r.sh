#!/usr/bin/env bash
# round 2, run R: pair kernels with the law as a compile-time fact: parity, dense vs tiled A/B on the compute-bound configurations
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_api2.py tests/test_hash.py tests/test_big_split.py -m gpu -x -q > $OUT/r2r_pytest.txt 2>&1; echo "pytest exit $?" >> $OUT/r2r_pytest.txt
tail -4 $OUT/r2r_pytest.txt
CHB_PAIR_TILE=1 timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > $OUT/r2r_pytest_tile.txt 2>&1; echo "pytest exit $?" >> $OUT/r2r_pytest_tile.txt
tail -3 $OUT/r2r_pytest_tile.txt
run() { local label="$1"; shift
  timeout 300 python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu --no-newton --no-configs --stages "$@" 2>$OUT/cfg_err.txt | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); s=d.get('stages') or {}
print('$label', 'ms', round(d['ms_per_step'],4), 'U1/s', round(d['value']), 'frac', round(d['roofline']['frac'],4), 'fusedver', d.get('fused_version'), 'launches', d['gpu_launches'], {k: round(v,4) for k,v in s.items()})" || tail -3 $OUT/cfg_err.txt
}
{
for knobs in "CHB_PAIR_TILE=0" "CHB_PAIR_TILE=1"; do
  env $knobs bash -c "$(declare -f run); OUT=$OUT; run 'C4 64C@16384 b256 [$knobs]' --system 64C --nint 16384 --batch 256"
  env $knobs bash -c "$(declare -f run); OUT=$OUT; run '16C@4096 b1024 [$knobs]' --system 16C --nint 4096 --batch 1024"
  env $knobs bash -c "$(declare -f run); OUT=$OUT; run '20B@1200 b1024 [$knobs]' --system 20B --nint 1200 --batch 1024"
done
run "C3 5Dq@10240 b1024" --system 5Dq --nint 10240 --batch 1024
run "C2 3B@4096 b4096" --system 3B --nint 4096 --batch 4096
} > $OUT/r2r_configs.txt 2>&1
cat $OUT/r2r_configs.txt
