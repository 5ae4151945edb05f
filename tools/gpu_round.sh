#!/usr/bin/env bash
# One GPU visit: parity tests, bench, ncu launch list.  Usage (under gpurun): bash tools/gpu_round.sh [tag]
TAG="${1:-r1}"
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/${TAG}_smi.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest exit $?" >> $OUT/${TAG}_pytest.log
tail -5 $OUT/${TAG}_pytest.log
timeout 600 python bench.py --steps 20 --warmup 5 --stages > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; echo "bench exit $?"
cat $OUT/${TAG}_bench.json
timeout 300 python bench.py --steps 2 --warmup 3 --batch 512 --no-e2e --no-cpu --no-newton > $OUT/${TAG}_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $OUT/${TAG}_launches.csv \
    python bench.py --steps 2 --warmup 3 --batch 512 --no-e2e --no-cpu --no-newton > $OUT/${TAG}_ncu.log 2>&1
echo "ncu exit $?"
if [ -n "$NCU_FULL" ]; then
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_fused -s 3 -c 1 -o $OUT/${TAG}_fused \
    python bench.py --steps 1 --warmup 3 --batch 592 --no-e2e --no-cpu --no-newton > $OUT/${TAG}_ncufull.log 2>&1
echo "ncu full exit $?"
fi
