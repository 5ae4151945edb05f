#!/usr/bin/env bash
# round 2, run C: new API tests on the GPU + v5 after the templated load/combine/store and shared-memory tables
OUT=gpurun_out; mkdir -p $OUT
timeout 1200 python -m pytest tests/test_api2.py tests/test_ode_forces.py tests/test_postsolve.py tests/test_hash.py -m gpu -x -q > $OUT/r2c_pytest.txt 2>&1; echo "pytest exit $?" >> $OUT/r2c_pytest.txt
tail -5 $OUT/r2c_pytest.txt
run() { local label="$1"; shift
  timeout 300 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu --no-newton --stages "$@" 2>$OUT/cfg_err.txt | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); s=d.get('stages') or {}
print('$label', 'ms', round(d['ms_per_step'],4), 'U1/s', round(d['value']), 'fusedver', d.get('fused_version'), 'launches', d['gpu_launches'], {k: round(v,4) for k,v in s.items()})" || tail -3 $OUT/cfg_err.txt
}
{
run "C3 5Dq@10240 b1024" --system 5Dq --nint 10240 --batch 1024
run "C3 5Dq@10240 b2072" --system 5Dq --nint 10240 --batch 2072
run "C3s 5Dq@1280 b4096" --system 5Dq --nint 1280 --batch 4096
run "3D@7680 b2048" --system 3D --nint 7680 --batch 2048
run "3C5k@15360 b2048" --system 3C5k --nint 15360 --batch 2048
} > $OUT/r2c_configs.txt 2>&1
cat $OUT/r2c_configs.txt
CMD="python bench.py --system 5Dq --nint 10240 --batch 296 --steps 1 --warmup 3 --no-e2e --no-cpu --no-newton"
$CMD > $OUT/r2c_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_fused5 -s 3 -c 1 -o /tmp/r2c_prof $CMD > $OUT/r2c_ncu.log 2>&1
ncu -i /tmp/r2c_prof.ncu-rep --page raw --csv > $OUT/r2c_raw.csv 2>/dev/null
ncu -i /tmp/r2c_prof.ncu-rep --page source --csv > $OUT/r2c_source.csv 2>/dev/null
tail -2 $OUT/r2c_ncu.log
