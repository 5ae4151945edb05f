#!/usr/bin/env bash
# round 2, run F: v4 after the conflict-free row accesses, uniform-weight pair kernels, drop-in test, sweep tool, default bench line
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests/test_dropin.py tests/test_gpu_parity.py tests/test_api2.py -m gpu -x -q > $OUT/r2f_pytest.txt 2>&1; echo "pytest exit $?" >> $OUT/r2f_pytest.txt
tail -4 $OUT/r2f_pytest.txt
run() { local label="$1"; shift
  timeout 300 python bench.py --steps 10 --warmup 3 --no-e2e --no-cpu --no-newton --no-configs --stages "$@" 2>$OUT/cfg_err.txt | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); s=d.get('stages') or {}
print('$label', 'ms', round(d['ms_per_step'],4), 'U1/s', round(d['value']), 'frac', round(d['roofline']['frac'],4), 'fusedver', d.get('fused_version'), 'launches', d['gpu_launches'], {k: round(v,4) for k,v in s.items()})" || tail -3 $OUT/cfg_err.txt
}
{
run "C2 3B@4096 b4096 v4" --system 3B --nint 4096 --batch 4096
run "C2 3B@4096 b4096 v4 (20 steps)" --system 3B --nint 4096 --batch 4096 --steps 20
run "C4 64C@16384 b256 dense uniform" --system 64C --nint 16384 --batch 256
CHB_PAIR_UNIFORM=0 bash -c "$(declare -f run); OUT=$OUT; run 'C4 64C@16384 b256 dense weights' --system 64C --nint 16384 --batch 256"
CHB_PAIR_TILE=1 bash -c "$(declare -f run); OUT=$OUT; run 'C4 64C@16384 b256 tile uniform' --system 64C --nint 16384 --batch 256"
run "20B@1200 b1024 dense uniform" --system 20B --nint 1200 --batch 1024
CHB_PAIR_TILE=1 bash -c "$(declare -f run); OUT=$OUT; run '20B@1200 b1024 tile' --system 20B --nint 1200 --batch 1024"
run "16C@4096 b1024" --system 16C --nint 4096 --batch 1024
CHB_PAIR_TILE=1 bash -c "$(declare -f run); OUT=$OUT; run '16C@4096 b1024 tile uniform' --system 16C --nint 4096 --batch 1024"
run "C3 5Dq@10240 b1024" --system 5Dq --nint 10240 --batch 1024
} > $OUT/r2f_configs.txt 2>&1
cat $OUT/r2f_configs.txt
timeout 600 python tools/sweep_bench.py --guesses 8192 --chunk 2048 --cpu-seeds 32 > $OUT/r2f_sweep_1gpu_8192.json 2> $OUT/r2f_sweep_err.txt; tail -c 1500 $OUT/r2f_sweep_1gpu_8192.json; tail -3 $OUT/r2f_sweep_err.txt
( time timeout 900 python bench.py ) > $OUT/r2f_bench_default.json 2> $OUT/r2f_bench_default_err.txt; tail -c 3000 $OUT/r2f_bench_default.json; tail -5 $OUT/r2f_bench_default_err.txt
( time timeout 600 python bench.py --impl reference --steps 3 --warmup 1 ) > $OUT/r2f_bench_reference.json 2> $OUT/r2f_bench_reference_err.txt; tail -c 1500 $OUT/r2f_bench_reference.json; tail -4 $OUT/r2f_bench_reference_err.txt
