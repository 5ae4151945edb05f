#!/usr/bin/env bash
# round 2, run D: full GPU suite; config table with the tiled pair kernel + parity-split transforms; launch list of 64C
OUT=gpurun_out; mkdir -p $OUT
timeout 1800 python -m pytest tests -m gpu -x -q > $OUT/r2d_pytest.txt 2>&1; echo "pytest exit $?" >> $OUT/r2d_pytest.txt
tail -5 $OUT/r2d_pytest.txt
run() { local label="$1"; shift
  timeout 300 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu --no-newton --stages "$@" 2>$OUT/cfg_err.txt | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); s=d.get('stages') or {}
print('$label', 'ms', round(d['ms_per_step'],4), 'U1/s', round(d['value']), 'fusedver', d.get('fused_version'), 'launches', d['gpu_launches'], {k: round(v,4) for k,v in s.items()})" || tail -3 $OUT/cfg_err.txt
}
{
run "C3 5Dq@10240 b1024" --system 5Dq --nint 10240 --batch 1024
run "C3s 5Dq@1280 b4096" --system 5Dq --nint 1280 --batch 4096
run "3D@7680 b2048" --system 3D --nint 7680 --batch 2048
run "3C5k@15360 b2048" --system 3C5k --nint 15360 --batch 2048
run "C4 64C@16384 b256" --system 64C --nint 16384 --batch 256
run "C4 64C@16384 b1024" --system 64C --nint 16384 --batch 1024
run "20B@1200 b1024" --system 20B --nint 1200 --batch 1024
run "16C@4096 b1024" --system 16C --nint 4096 --batch 1024
CHB_PAIR_TILE=0 CHB_BIG=0 bash -c "$(declare -f run); OUT=$OUT; run 'C4 64C@16384 b256 old kernels' --system 64C --nint 16384 --batch 256"
CHB_PAIR_TILE=0 bash -c "$(declare -f run); OUT=$OUT; run 'C4 64C@16384 b256 dense pair + split transforms' --system 64C --nint 16384 --batch 256"
} > $OUT/r2d_configs.txt 2>&1
cat $OUT/r2d_configs.txt
CMD="python bench.py --system 64C --nint 16384 --batch 148 --steps 1 --warmup 3 --no-e2e --no-cpu --no-newton"
$CMD > $OUT/r2d_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"k_big|k_pair_tile" -s 9 -c 3 -o /tmp/r2d_prof $CMD > $OUT/r2d_ncu.log 2>&1
ncu -i /tmp/r2d_prof.ncu-rep --page raw --csv > $OUT/r2d_raw.csv 2>/dev/null
tail -2 $OUT/r2d_ncu.log
CMD="python bench.py --system 5Dq --nint 10240 --batch 296 --steps 1 --warmup 3 --no-e2e --no-cpu --no-newton"
$CMD > $OUT/r2d_plain5.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_fused5 -s 3 -c 1 -o /tmp/r2d_prof5 $CMD > $OUT/r2d_ncu5.log 2>&1
ncu -i /tmp/r2d_prof5.ncu-rep --page raw --csv > $OUT/r2d_raw5.csv 2>/dev/null
ncu -i /tmp/r2d_prof5.ncu-rep --page source --csv > $OUT/r2d_source5.csv 2>/dev/null
