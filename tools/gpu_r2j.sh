#!/usr/bin/env bash
# round 2, run J: where a Newton iteration at a deep refinement level spends its time (launch list by kernel), sweep with worker threads
OUT=gpurun_out; mkdir -p $OUT
timeout 300 python tools/newton_deep_probe.py --reps 2 > $OUT/r2j_probe.txt 2>&1; cat $OUT/r2j_probe.txt
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file /tmp/r2j_launches.csv python tools/newton_deep_probe.py --reps 1 > $OUT/r2j_ncu_stdout.txt 2>&1
python - <<'PY' > gpurun_out/r2j_launch_summary.txt 2>&1
import csv, collections, re
rows = []
with open('/tmp/r2j_launches.csv') as fh:
    lines = [l for l in fh if not l.startswith('==')]
rd = csv.DictReader(lines)
agg = collections.OrderedDict()
n = 0
for r in rd:
    if r.get('Metric Name') != 'gpu__time_duration.sum': continue
    name = re.sub(r'\(.*', '', r['Kernel Name'])
    v = float(r['Metric Value'].replace(',', ''))
    unit = r['Metric Unit']
    if unit == 'ns': v /= 1e3
    elif unit == 'ms': v *= 1e3
    a = agg.setdefault(name, [0, 0.0]); a[0] += 1; a[1] += v; n += 1
tot = sum(a[1] for a in agg.values())
print('launches', n, 'total us', round(tot, 1))
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f'{a[1]:12.1f} us {100 * a[1] / tot:5.1f} %  {a[0]:6d} launches  {a[1] / a[0]:9.1f} us each  {k}')
PY
cat gpurun_out/r2j_launch_summary.txt | head -40
for w in 4 8; do
timeout 900 python tools/sweep_bench.py --guesses 8192 --chunk 1024 --workers $w > $OUT/r2j_sweep_8192_w$w.json 2> $OUT/r2j_sweep_w${w}_err.txt; head -c 700 $OUT/r2j_sweep_8192_w$w.json; echo; tail -3 $OUT/r2j_sweep_w${w}_err.txt
done
