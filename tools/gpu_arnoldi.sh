#!/usr/bin/env bash
# Arnoldi / GMRES-step kernels: parity, solves/hour with each level (CHB_NO_GMRES_STEP=1: Arnoldi kernel + array-level Givens;
# CHB_NO_ARNOLDI=1: array-level Gram-Schmidt), torch profile of the driver.
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests/test_arnoldi.py tests/test_gpu_parity.py -m gpu -x -q -k "arnoldi or newton" > $OUT/arn_pytest.log 2>&1; echo "pytest exit $?"; tail -3 $OUT/arn_pytest.log
timeout 900 python tools/solve_bench.py --batch 8192 --ref-seeds 4 > $OUT/arn_solve_8192.json 2> $OUT/arn_solve_8192.err; echo "solve exit $?"; tail -2 $OUT/arn_solve_8192.json
CHB_NO_GMRES_STEP=1 timeout 900 python tools/solve_bench.py --batch 8192 --ref-seeds 0 > $OUT/arn_solve_8192_nostep.json 2> $OUT/arn_solve_8192_nostep.err; echo "solve(no step) exit $?"; tail -2 $OUT/arn_solve_8192_nostep.json
CHB_NO_ARNOLDI=1 timeout 900 python tools/solve_bench.py --batch 8192 --ref-seeds 0 > $OUT/arn_solve_8192_einsum.json 2> $OUT/arn_solve_8192_einsum.err; echo "solve(einsum) exit $?"; tail -2 $OUT/arn_solve_8192_einsum.json
timeout 900 python tools/solve_bench.py --batch 16384 --ref-seeds 0 > $OUT/arn_solve_16384.json 2> $OUT/arn_solve_16384.err; echo "solve 16k exit $?"; tail -2 $OUT/arn_solve_16384.json
timeout 900 python tools/newton_profile.py > $OUT/arn_newton_profile.txt 2>&1; echo "profile exit $?"; head -40 $OUT/arn_newton_profile.txt
