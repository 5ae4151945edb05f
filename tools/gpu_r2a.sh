#!/usr/bin/env bash
# round 2, run A: GPU parity suite + config table with the general fused kernel (k_fused5)
OUT=gpurun_out; mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > $OUT/r2a_gpu.txt
timeout 1500 python -m pytest tests -m gpu -x -q > $OUT/r2a_pytest.txt 2>&1; echo "pytest exit $?" >> $OUT/r2a_pytest.txt
tail -5 $OUT/r2a_pytest.txt
bash tools/gpu_configs.sh > $OUT/r2a_configs.txt 2>&1
run() { local label="$1"; shift
  timeout 300 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu --no-newton --stages "$@" 2>$OUT/cfg_err.txt | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); s=d.get('stages') or {}
print('$label', 'ms', round(d['ms_per_step'],4), 'U1/s', round(d['value']), 'fusedver', d.get('fused_version'), 'launches', d['gpu_launches'], {k: round(v,4) for k,v in s.items()})" || tail -3 $OUT/cfg_err.txt
}
run "3D@7680 b2048" --system 3D --nint 7680 --batch 2048 >> $OUT/r2a_configs.txt 2>&1
run "3C5k@15360 b2048" --system 3C5k --nint 15360 --batch 2048 >> $OUT/r2a_configs.txt 2>&1
CHB_FUSED5=0 bash -c "$(declare -f run); OUT=$OUT; run 'C3 5Dq@10240 b1024 staged(CHB_FUSED5=0)' --system 5Dq --nint 10240 --batch 1024" >> $OUT/r2a_configs.txt 2>&1
cat $OUT/r2a_configs.txt
