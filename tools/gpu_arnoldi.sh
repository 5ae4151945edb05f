#!/usr/bin/env bash
# GMRES kernels of the Newton driver: parity, micro-benchmark of the step kernel, solves/hour, torch profile of the driver.
#   CHB_ARNOLDI_RING=0      register-prefetch Arnoldi kernel instead of the bulk-copy ring
#   CHB_GMRES_TAIL_SPLIT=0  Givens tail inside the step kernel;  CHB_NO_GMRES_LIB=1 / CHB_NO_GMRES_STEP=1 / CHB_NO_ARNOLDI=1: older levels
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests/test_arnoldi.py tests/test_gpu_parity.py -m gpu -x -q -k "arnoldi or newton or masked or gmres or find" > $OUT/arn_pytest.log 2>&1; echo "pytest exit $?"; tail -3 $OUT/arn_pytest.log
timeout 600 python tools/arnoldi_bench.py --batches 1024,8192 > $OUT/arnoldi_bench.txt 2>&1; grep "arnoldi " $OUT/arnoldi_bench.txt
for B in 8192 16384 32768; do
timeout 900 python tools/solve_bench.py --batch $B --ref-seeds 4 > $OUT/ring_solve_$B.json 2> $OUT/ring_solve_$B.err; echo "solve $B exit $?"; tail -1 $OUT/ring_solve_$B.json | cut -c1-420
done
timeout 900 python tools/newton_profile.py > $OUT/ring_newton_profile.txt 2>&1; echo "profile exit $?"; grep "k_arnoldi\|k_fused\|k_gmres\|k_givens\|Self CUDA time total\|Self CPU time total" $OUT/ring_newton_profile.txt | cut -c1-60,120-230
