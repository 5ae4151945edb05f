#!/usr/bin/env bash
# round 2, run B: ncu --set full of k_fused5 on 5Dq nint 10240 (2 states per persistent CTA); the report is exported to CSV
# on the box (it embeds the whole fatbin and exceeds the size gpurun copies back)
OUT=gpurun_out; mkdir -p $OUT
CMD="python bench.py --system 5Dq --nint 10240 --batch 296 --steps 1 --warmup 3 --no-e2e --no-cpu --no-newton"
$CMD > $OUT/r2b_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_fused5 -s 3 -c 1 -o /tmp/r2b_fused5_5Dq $CMD > $OUT/r2b_ncu.log 2>&1
ncu -i /tmp/r2b_fused5_5Dq.ncu-rep --page raw --csv > $OUT/r2b_raw.csv 2>/dev/null
ncu -i /tmp/r2b_fused5_5Dq.ncu-rep --page details --csv > $OUT/r2b_details.csv 2>/dev/null
ncu -i /tmp/r2b_fused5_5Dq.ncu-rep --page source --csv > $OUT/r2b_source.csv 2>/dev/null
ls -la /tmp/r2b_fused5_5Dq.ncu-rep $OUT
tail -3 $OUT/r2b_plain.log; tail -5 $OUT/r2b_ncu.log
