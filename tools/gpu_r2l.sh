#!/usr/bin/env bash
# round 2, run L: staged kernels leave out stopped states; launch profile of a Newton iteration on the fused path with few live states; sweeps
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests/test_arnoldi.py tests/test_newton.py tests/test_refine.py tests/test_big_split.py -m gpu -x -q > $OUT/r2l_pytest.txt 2>&1; echo "pytest exit $?" >> $OUT/r2l_pytest.txt
tail -4 $OUT/r2l_pytest.txt
timeout 300 python tools/newton_deep_probe.py --reps 2 > $OUT/r2l_probe.txt 2>&1; cat $OUT/r2l_probe.txt
timeout 300 python tools/newton_deep_probe.py --reps 2 --nint 3072 --states 8 --maxiter 20 > $OUT/r2l_probe_3072.txt 2>&1; cat $OUT/r2l_probe_3072.txt
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file /tmp/r2l_launches.csv python tools/newton_deep_probe.py --reps 1 --nint 3072 --states 8 --maxiter 20 > $OUT/r2l_ncu_stdout.txt 2>&1
python - <<'PY' > gpurun_out/r2l_launch_summary_3072.txt 2>&1
import csv, collections, re
with open('/tmp/r2l_launches.csv') as fh:
    lines = [l for l in fh if not l.startswith('==')]
agg = collections.OrderedDict(); n = 0
for r in csv.DictReader(lines):
    if r.get('Metric Name') != 'gpu__time_duration.sum': continue
    name = re.sub(r'\(.*', '', r['Kernel Name'])
    v = float(r['Metric Value'].replace(',', '')); unit = r['Metric Unit']
    if unit == 'ns': v /= 1e3
    elif unit == 'ms': v *= 1e3
    a = agg.setdefault(name, [0, 0.0]); a[0] += 1; a[1] += v; n += 1
tot = sum(a[1] for a in agg.values())
print('launches', n, 'total us', round(tot, 1))
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f'{a[1]:12.1f} us {100 * a[1] / tot:5.1f} %  {a[0]:6d} launches  {a[1] / a[0]:9.1f} us each  {k}')
PY
head -24 gpurun_out/r2l_launch_summary_3072.txt
timeout 900 python tools/sweep_bench.py --guesses 8192 --chunk 8192 --trace $OUT/r2l_trace_8192.jsonl > $OUT/r2l_sweep_8192_onechunk.json 2> $OUT/r2l_sweep_8192_err.txt; head -c 600 $OUT/r2l_sweep_8192_onechunk.json; echo; tail -3 $OUT/r2l_sweep_8192_err.txt
timeout 900 python tools/sweep_bench.py --guesses 8192 --chunk 1024 --workers 4 > $OUT/r2l_sweep_8192_w4.json 2> $OUT/r2l_sweep_w4_err.txt; head -c 600 $OUT/r2l_sweep_8192_w4.json; echo; tail -3 $OUT/r2l_sweep_w4_err.txt
timeout 900 python tools/sweep_bench.py --guesses 32768 --chunk 4096 --workers 8 --cpu-seeds 32 > $OUT/r2l_sweep_32768_w8.json 2> $OUT/r2l_sweep_32768_err.txt; head -c 600 $OUT/r2l_sweep_32768_w8.json; echo; tail -3 $OUT/r2l_sweep_32768_err.txt
