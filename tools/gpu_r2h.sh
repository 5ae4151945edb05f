#!/usr/bin/env bash
# round 2, run H: where the sweep spends its time (trace per (refinement level, optimiser level) group), bounded default bench, reference arm, GPU tests
OUT=gpurun_out; mkdir -p $OUT
timeout 600 python tools/sweep_bench.py --guesses 2048 --chunk 2048 --trace $OUT/r2h_trace_2048.jsonl > $OUT/r2h_sweep_2048.json 2> $OUT/r2h_sweep_2048_err.txt; tail -c 600 $OUT/r2h_sweep_2048.json
timeout 900 python tools/sweep_bench.py --guesses 8192 --chunk 8192 --trace $OUT/r2h_trace_8192.jsonl > $OUT/r2h_sweep_8192_onechunk.json 2> $OUT/r2h_sweep_8192_err.txt; tail -c 600 $OUT/r2h_sweep_8192_onechunk.json
( time timeout 900 python bench.py ) > $OUT/r2h_bench_default.json 2> $OUT/r2h_bench_default_err.txt; tail -c 3000 $OUT/r2h_bench_default.json; tail -5 $OUT/r2h_bench_default_err.txt
( time timeout 900 python bench.py --impl reference --steps 3 --warmup 1 ) > $OUT/r2h_bench_reference.json 2> $OUT/r2h_bench_reference_err.txt; tail -c 1500 $OUT/r2h_bench_reference.json; tail -4 $OUT/r2h_bench_reference_err.txt
timeout 1500 python -m pytest tests -m gpu -x -q > $OUT/r2h_pytest_gpu.txt 2>&1; tail -5 $OUT/r2h_pytest_gpu.txt
