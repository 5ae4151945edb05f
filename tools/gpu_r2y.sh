#!/usr/bin/env bash
# round 2, run Y: Newton iteration of a few states at nint 3072 / 6144: fused kernel against the staged kernels (latency of one state)
OUT=gpurun_out; mkdir -p $OUT
{
for st in 1 8 32; do
  for ni in 3072 6144 12288; do
    timeout 120 python tools/newton_deep_probe.py --reps 1 --nint $ni --states $st --maxiter 10 | tail -1
    timeout 120 python tools/newton_deep_probe.py --reps 1 --nint $ni --states $st --maxiter 10 --no-fused | tail -1
    timeout 120 python tools/newton_deep_probe.py --reps 1 --nint $ni --states $st --maxiter 10 --no-fused --split 2 | tail -1
  done
done
} > $OUT/r2y_probe.txt 2>&1
cat $OUT/r2y_probe.txt
