#!/usr/bin/env bash
# BASELINE configs other than the bench line (parity cases): throughput of U1 and of the staged operators, for profiles/
OUT=gpurun_out; mkdir -p $OUT
run() { local label="$1"; shift
  timeout 300 python bench.py --steps 5 --warmup 3 --no-e2e --no-cpu --stages "$@" 2>$OUT/cfg_err.txt | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); s=d.get('stages') or {}
print('$label', 'ms', round(d['ms_per_step'],4), 'U1/s', round(d['value']), 'fusedver', d.get('fused_version'), 'launches', d['gpu_launches'], {k: round(v,4) for k,v in s.items()})" || tail -3 $OUT/cfg_err.txt
}
run "C1 3C@3072 b4096" --system 3C --nint 3072 --batch 4096
run "C2b 3C@12288 b4096" --system 3C --nint 12288 --batch 4096
run "C3 5Dq@10240 b1024" --system 5Dq --nint 10240 --batch 1024
run "C3s 5Dq@1280 b1024" --system 5Dq --nint 1280 --batch 1024
run "C4 64C@16384 b256" --system 64C --nint 16384 --batch 256
run "20B@1200 b1024" --system 20B --nint 1200 --batch 1024
