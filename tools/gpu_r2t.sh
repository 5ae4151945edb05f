#!/usr/bin/env bash
# round 2, run T: ncu --set full of the tiled pair kernel (64C, 148 states)
OUT=gpurun_out; mkdir -p $OUT
CMD="python bench.py --system 64C --nint 16384 --batch 148 --steps 1 --warmup 3 --no-e2e --no-cpu --no-newton --no-configs"
$CMD > $OUT/r2t_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"k_pair_tile" -s 3 -c 1 -o /tmp/r2t_prof $CMD > $OUT/r2t_ncu.log 2>&1
tail -2 $OUT/r2t_ncu.log
ncu -i /tmp/r2t_prof.ncu-rep --page raw --csv > $OUT/r2t_raw.csv 2>/dev/null
ncu -i /tmp/r2t_prof.ncu-rep --page source --csv > $OUT/r2t_source_sass.csv 2>/dev/null
ls -la $OUT/r2t_*
