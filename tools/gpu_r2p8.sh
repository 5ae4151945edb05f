#!/usr/bin/env bash
# round 2, run P (8 GPUs): host-link ceiling with 8 concurrent ranks (with / without NUMA pinning), bench --gpus 8 (U1 + host-buffer leg), sweep on 8 GPUs
OUT=gpurun_out; mkdir -p $OUT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511"
nvidia-smi topo -m > $OUT/r2p_topo.txt 2>&1
timeout 200 $TR tools/pcie_ranks.py --iters 6 > $OUT/r2p_pcie_8.json 2> $OUT/r2p_pcie_8_err.txt; tail -1 $OUT/r2p_pcie_8.json
timeout 200 $TR tools/pcie_ranks.py --iters 6 --pin-numa > $OUT/r2p_pcie_8_pinned.json 2> $OUT/r2p_pcie_8_pinned_err.txt; tail -1 $OUT/r2p_pcie_8_pinned.json
timeout 100 python tools/pcie_ranks.py --iters 6 > $OUT/r2p_pcie_1.json 2>/dev/null; tail -1 $OUT/r2p_pcie_1.json
timeout 400 $TR bench.py --gpus 8 --steps 10 --warmup 3 --no-newton --no-configs > $OUT/r2p_bench_8gpu_pinned.json 2> $OUT/r2p_bench_8gpu_pinned_err.txt; tail -c 900 $OUT/r2p_bench_8gpu_pinned.json
timeout 400 $TR bench.py --gpus 8 --steps 10 --warmup 3 --no-newton --no-configs --no-pin-numa > $OUT/r2p_bench_8gpu_unpinned.json 2> $OUT/r2p_bench_8gpu_unpinned_err.txt; tail -c 900 $OUT/r2p_bench_8gpu_unpinned.json
timeout 600 $TR tools/sweep_bench.py --guesses 131072 --chunk 4096 --workers 4 > $OUT/r2p_sweep_8gpu_131072.json 2> $OUT/r2p_sweep_8gpu_err.txt; head -c 700 $OUT/r2p_sweep_8gpu_131072.json; tail -3 $OUT/r2p_sweep_8gpu_err.txt
timeout 400 $TR bench.py --gpus 8 --steps 10 --warmup 3 --no-e2e --no-configs --newton-reps 1 > $OUT/r2p_bench_8gpu_newton.json 2> $OUT/r2p_bench_8gpu_newton_err.txt; tail -c 600 $OUT/r2p_bench_8gpu_newton.json
