#!/usr/bin/env python
"""Host-link ceiling of the e2e leg with N concurrent ranks: every rank copies `--mb` MB of pinned host memory
to its GPU and the same amount back, both directions at once on two streams (what chb_run_host does per step: parameters up,
gradients down), all ranks between barriers.  Prints the aggregate GB/s per direction and the U1/s ceiling it implies for the
bench workload (3B nint 4096: 196 560 B up + 196 568 B down per state).

    python tools/pcie_ranks.py                                                   # 1 rank
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 tools/pcie_ranks.py [--pin-numa]
`--pin-numa`: bind the process (and so the first touch of its pinned buffers) to the CPU set nvidia-smi reports for its GPU."""
import argparse
import json
import os
import subprocess
import time


def gpu_cpu_affinity(index):
    try:
        out = subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True, timeout=20).stdout
        for line in out.splitlines():
            f = line.split()
            if f and f[0] == f"GPU{index}":
                for tok in f[1:]:
                    if tok[0].isdigit() and ("-" in tok or "," in tok) and not tok.endswith("X"):
                        cpus = set()
                        for part in tok.split(","):
                            a, _, b = part.partition("-")
                            cpus.update(range(int(a), int(b or a) + 1))
                        return cpus
    except Exception:
        pass
    return None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mb", type=int, default=805)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--pin-numa", action="store_true")
    a = ap.parse_args()
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    aff = gpu_cpu_affinity(local)
    if a.pin_numa and aff:
        os.sched_setaffinity(0, aff)
    import torch
    import torch.distributed as dist
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    n = a.mb * 1000 * 1000 // 8
    h_in = torch.empty(n, dtype=torch.float64).pin_memory()
    h_out = torch.empty(n, dtype=torch.float64).pin_memory()
    h_in.fill_(1.0)
    d_in = torch.empty(n, dtype=torch.float64, device=dev)
    d_out = torch.ones(n, dtype=torch.float64, device=dev)
    s_up, s_dn = torch.cuda.Stream(), torch.cuda.Stream()

    def step():
        with torch.cuda.stream(s_up):
            d_in.copy_(h_in, non_blocking=True)
        with torch.cuda.stream(s_dn):
            h_out.copy_(d_out, non_blocking=True)

    for _ in range(2):
        step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(a.iters):
        step()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([dt], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt = float(t.item())
    if rank == 0:
        gbs = world * a.iters * n * 8 / dt / 1e9
        print(json.dumps({"ranks": world, "mb_per_direction_per_step": a.mb, "pin_numa": bool(a.pin_numa), "cpu_affinity_gpu0": sorted(aff)[:4] + ["..."] + sorted(aff)[-1:] if aff else None,
                          "aggregate_gbs_per_direction": gbs, "per_rank_gbs_per_direction": gbs / world,
                          "u1_per_s_ceiling_3B_4096": gbs * 1e9 / 196568.0, "seconds": dt}))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
