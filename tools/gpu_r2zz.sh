#!/usr/bin/env bash
# round 2, last run: whole GPU suite and the default bench line on the final library
OUT=gpurun_out; mkdir -p $OUT
timeout 1500 python -m pytest tests -m gpu -x -q > $OUT/r2zz_pytest_gpu.txt 2>&1; echo "pytest exit $?" >> $OUT/r2zz_pytest_gpu.txt
tail -4 $OUT/r2zz_pytest_gpu.txt
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $OUT/r2zz_smoke.txt 2>&1; tail -2 $OUT/r2zz_smoke.txt
( time timeout 600 python bench.py ) > $OUT/r2zz_bench_default.json 2> $OUT/r2zz_bench_default_err.txt; tail -c 1200 $OUT/r2zz_bench_default.json; tail -4 $OUT/r2zz_bench_default_err.txt
