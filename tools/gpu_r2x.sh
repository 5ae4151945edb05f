#!/usr/bin/env bash
# round 2, run X: fewer launches per Newton iteration: parity tests, launch counts of the probes, search runner at the bench leg's settings
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests/test_newton.py tests/test_refine.py tests/test_dropin.py tests/test_arnoldi.py -m gpu -x -q > $OUT/r2x_pytest.txt 2>&1; echo "pytest exit $?" >> $OUT/r2x_pytest.txt
tail -3 $OUT/r2x_pytest.txt
timeout 300 python tools/newton_deep_probe.py --reps 2 --nint 3072 --states 8 --maxiter 20 > $OUT/r2x_probe_3072.txt 2>&1; cat $OUT/r2x_probe_3072.txt
timeout 300 python tools/newton_deep_probe.py --reps 2 > $OUT/r2x_probe_49152.txt 2>&1; cat $OUT/r2x_probe_49152.txt
sw() { local tag="$1"; shift
  timeout 400 python tools/sweep_bench.py "$@" > $OUT/r2x_sweep_$tag.json 2> $OUT/r2x_sweep_${tag}_err.txt
  python -c "
import json
d=json.loads([l for l in open('$OUT/r2x_sweep_$tag.json') if l.startswith('{')][-1])
print('$tag', {k:d.get(k) for k in ['guesses','chunk','workers_per_gpu','seconds','solves_per_hour','found','collision','not_converged']})" || tail -3 $OUT/r2x_sweep_${tag}_err.txt
}
{
sw c1024_w4 --guesses 8192 --chunk 1024 --workers 4 --reconverge 1
sw c1024_w8 --guesses 8192 --chunk 1024 --workers 8 --reconverge 1
sw full_c1024_w4 --guesses 8192 --chunk 1024 --workers 4
} > $OUT/r2x_summary.txt 2>&1
cat $OUT/r2x_summary.txt
