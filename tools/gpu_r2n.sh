#!/usr/bin/env bash
# round 2, run N: ncu --set full of the residue-split transforms on 64C (where do the cycles of a half-length FFT CTA go)
OUT=gpurun_out; mkdir -p $OUT
CMD="python bench.py --system 64C --nint 16384 --batch 148 --steps 1 --warmup 3 --no-e2e --no-cpu --no-newton --no-configs --stages"
$CMD > $OUT/r2n_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"k_big" -s 6 -c 2 -o /tmp/r2n_prof $CMD > $OUT/r2n_ncu.log 2>&1
tail -2 $OUT/r2n_ncu.log
ncu -i /tmp/r2n_prof.ncu-rep --page raw --csv > $OUT/r2n_raw.csv 2>/dev/null
ncu -i /tmp/r2n_prof.ncu-rep --page source --csv > $OUT/r2n_source_sass.csv 2>/dev/null
ncu -i /tmp/r2n_prof.ncu-rep --page source --print-source cuda --csv > $OUT/r2n_source_cuda.csv 2>$OUT/r2n_source_cuda_err.txt
ls -la $OUT/r2n_*; head -c 300 $OUT/r2n_source_cuda.csv
