#!/usr/bin/env bash
# round 2: ncu launch list of the bench command (U1 leg + e2e leg) and one --set full capture of the headline kernel
OUT=gpurun_out; mkdir -p $OUT
CMD="python bench.py --steps 2 --warmup 3 --no-newton --no-configs --no-cpu"
$CMD > $OUT/r2ncu_plain.json 2> $OUT/r2ncu_plain_err.txt && \
timeout 240 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/r2ncu_launches.csv $CMD > $OUT/r2ncu_stdout.txt 2>&1
CMD2="python bench.py --steps 1 --warmup 3 --no-newton --no-configs --no-cpu --no-e2e"
timeout 200 ncu --set full --clock-control none --import-source on -k regex:k_fused4 -s 3 -c 1 -o /tmp/r2ncu_full $CMD2 > $OUT/r2ncu_full_stdout.txt 2>&1
ncu -i /tmp/r2ncu_full.ncu-rep --page raw --csv > $OUT/r2ncu_full_raw.csv 2>/dev/null
ls -la $OUT/r2ncu_*
