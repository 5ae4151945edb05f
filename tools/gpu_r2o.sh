#!/usr/bin/env bash
# round 2, run O: the no-allocation Hessian-vector entry point alone in a fresh process, then the whole GPU suite (it runs last), default bench, reference arm
OUT=gpurun_out; mkdir -p $OUT
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > $OUT/r2o_pytest_noalloc_alone.txt 2>&1; echo "pytest exit $?" >> $OUT/r2o_pytest_noalloc_alone.txt
tail -6 $OUT/r2o_pytest_noalloc_alone.txt
timeout 1800 python -m pytest tests -m gpu -x -q > $OUT/r2o_pytest_gpu.txt 2>&1; echo "pytest exit $?" >> $OUT/r2o_pytest_gpu.txt
tail -6 $OUT/r2o_pytest_gpu.txt
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $OUT/r2o_smoke.txt 2>&1; tail -2 $OUT/r2o_smoke.txt
( time timeout 900 python bench.py ) > $OUT/r2o_bench_default.json 2> $OUT/r2o_bench_default_err.txt; tail -c 1500 $OUT/r2o_bench_default.json; tail -5 $OUT/r2o_bench_default_err.txt
( time timeout 900 python bench.py --impl reference --steps 3 --warmup 1 ) > $OUT/r2o_bench_reference.json 2> $OUT/r2o_bench_reference_err.txt; tail -c 800 $OUT/r2o_bench_reference.json; tail -4 $OUT/r2o_bench_reference_err.txt
