#!/usr/bin/env bash
# GMRES kernels of the Newton driver: parity, solves/hour at each level, torch profile of the driver.
#   CHB_NO_GMRES_LIB=1   gather/scatter of the active rows + array-level back substitution (gmres_step kernel kept)
#   CHB_NO_GMRES_STEP=1  Arnoldi kernel + array-level Givens;   CHB_NO_ARNOLDI=1  array-level Gram-Schmidt
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests/test_arnoldi.py tests/test_gpu_parity.py -m gpu -x -q -k "arnoldi or newton or masked or gmres or fused" > $OUT/arn_pytest.log 2>&1; echo "pytest exit $?"; tail -3 $OUT/arn_pytest.log
for B in 8192 16384 32768; do
timeout 900 python tools/solve_bench.py --batch $B --ref-seeds 4 > $OUT/lib_solve_$B.json 2> $OUT/lib_solve_$B.err; echo "solve $B exit $?"; tail -1 $OUT/lib_solve_$B.json | cut -c1-700
done
CHB_NO_GMRES_LIB=1 timeout 900 python tools/solve_bench.py --batch 8192 --ref-seeds 0 > $OUT/lib_solve_8192_nolib.json 2> $OUT/lib_solve_8192_nolib.err; echo "solve(no lib) exit $?"; tail -1 $OUT/lib_solve_8192_nolib.json | cut -c1-500
timeout 900 python tools/newton_profile.py > $OUT/lib_newton_profile.txt 2>&1; echo "profile exit $?"; head -36 $OUT/lib_newton_profile.txt
