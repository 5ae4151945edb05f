#!/usr/bin/env bash
# round 2, run S: tiled pair kernel as the default: full GPU suite, launch list of the staged 64C step, config lines, default bench
OUT=gpurun_out; mkdir -p $OUT
timeout 1800 python -m pytest tests -m gpu -x -q > $OUT/r2s_pytest_gpu.txt 2>&1; echo "pytest exit $?" >> $OUT/r2s_pytest_gpu.txt
tail -4 $OUT/r2s_pytest_gpu.txt
CMD="python bench.py --system 64C --nint 16384 --batch 256 --steps 2 --warmup 3 --no-e2e --no-cpu --no-newton --no-configs"
$CMD > $OUT/r2s_plain.log 2>&1 && timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file $OUT/r2s_launches_64C_b256.csv $CMD > $OUT/r2s_ncu.log 2>&1
python - <<'PY'
import csv
rows=[l for l in open('gpurun_out/r2s_launches_64C_b256.csv') if not l.startswith('==')]
for r in list(csv.DictReader(rows))[-8:]:
    print(r['Kernel Name'][:60], r['Metric Value'], r['Metric Unit'])
PY
( time timeout 900 python bench.py ) > $OUT/r2s_bench_default.json 2> $OUT/r2s_bench_default_err.txt; tail -c 2600 $OUT/r2s_bench_default.json; tail -4 $OUT/r2s_bench_default_err.txt
