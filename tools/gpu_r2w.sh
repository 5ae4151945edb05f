#!/usr/bin/env bash
# round 2, run W: chunk size / chunks in flight of the search runner at the bench leg's settings (8192 starts, n_reconverge_it_max 1)
OUT=gpurun_out; mkdir -p $OUT
sw() { local tag="$1"; shift
  timeout 400 python tools/sweep_bench.py "$@" > $OUT/r2w_sweep_$tag.json 2> $OUT/r2w_sweep_${tag}_err.txt
  python -c "
import json
d=json.loads([l for l in open('$OUT/r2w_sweep_$tag.json') if l.startswith('{')][-1])
print('$tag', {k:d.get(k) for k in ['guesses','chunk','workers_per_gpu','seconds','solves_per_hour','found','collision','not_converged']})" || tail -3 $OUT/r2w_sweep_${tag}_err.txt
}
{
sw c1024_w4 --guesses 8192 --chunk 1024 --workers 4 --reconverge 1
sw c512_w8 --guesses 8192 --chunk 512 --workers 8 --reconverge 1
sw c512_w16 --guesses 8192 --chunk 512 --workers 16 --reconverge 1
sw c256_w16 --guesses 8192 --chunk 256 --workers 16 --reconverge 1
sw c1024_w8 --guesses 8192 --chunk 1024 --workers 8 --reconverge 1
} > $OUT/r2w_summary.txt 2>&1
cat $OUT/r2w_summary.txt
