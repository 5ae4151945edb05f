#!/usr/bin/env python
"""Sharded orbit-search sweep (BASELINE config 5): N independent initial guesses split contiguously over the ranks (one process
per GPU, plan replicated, NO data-path collective), each rank runs the batched Newton-Krylov driver on its shard in chunks,
rank 0 gathers (seed, status, action, |grad|, iterations) host-side and prints one JSON line with solves/hour.

    python tools/sweep_bench.py --guesses 4096                                        # 1 GPU
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 tools/sweep_bench.py --guesses 32768
Starts: x = x_min + (x_max - x_min) * U[0,1) with np.random.seed(seed) (find.py:105-108,158), bounds from
PlanTables.make_params_bounds(nint, 0.1, 3, 50, 1e-16).
"""
import argparse, json, os, sys, time
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--system", default="3C"); ap.add_argument("--nint", type=int, default=3072)
    ap.add_argument("--guesses", type=int, default=4096); ap.add_argument("--chunk", type=int, default=4096)
    ap.add_argument("--f-tol", type=float, default=1e-9); ap.add_argument("--maxiter", type=int, default=60)
    a = ap.parse_args()
    import torch
    import torch.distributed as dist
    from choreographies2_b200 import NBodySystB200, PlanTables, sweep
    from choreographies2_b200.newton import newton_krylov_batched
    rank, world, local = sweep.env_rank_world()
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    T = PlanTables.load(os.path.join(ROOT, "choreographies2_b200", "systems", f"{a.system}.npz"))
    nbs = NBodySystB200(T, a.nint, device=local)
    x_min, x_max = T.make_params_bounds(a.nint, 0.1, 3, 50, 1e-16)
    lo, hi = sweep.shard_bounds(a.guesses, rank, world)

    def starts(s0, s1):
        out = np.empty((s1 - s0, nbs.nparams))
        for i, seed in enumerate(range(s0, s1)):
            np.random.seed(seed)
            out[i] = x_min + (x_max - x_min) * np.random.random(nbs.nparams)
        return out

    newton_krylov_batched(nbs, torch.from_numpy(starts(lo, min(hi, lo + 8))).cuda(), f_tol=a.f_tol, maxiter=2)  # warm-up
    x0_chunks = [(s0, min(hi, s0 + a.chunk)) for s0 in range(lo, hi, a.chunk)]
    x0_dev = [torch.from_numpy(starts(s0, s1)).cuda() for s0, s1 in x0_chunks]   # guesses resident before the timed region
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    rec = {"status": [], "action": [], "gradnorm": [], "nit": []}
    nU1 = nU2 = 0
    for x0 in x0_dev:
        r = newton_krylov_batched(nbs, x0, f_tol=a.f_tol, maxiter=a.maxiter)
        rec["status"].append(r.status); rec["nit"].append(r.nit); rec["gradnorm"].append(r.fnorm)
        rec["action"].append(nbs.params_to_action(r.x).cpu().numpy())
        nU1 += r.n_grad_evals; nU2 += r.n_hess_evals
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([dt, float(nU1), float(nU2)], dtype=torch.float64, device="cuda")
        tmax = t.clone(); dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        dt, nU1, nU2 = float(tmax[0]), int(t[1]), int(t[2])
    local_rec = {k: np.concatenate(v) if v else np.zeros((0,)) for k, v in rec.items()}
    merged = sweep.gather_records(local_rec, np.arange(lo, hi, dtype=np.int64))
    if rank == 0:
        st = merged["status"]
        conv = st == 1
        acts = np.round(merged["action"][conv], 6)
        print(json.dumps({
            "metric": "Newton orbit solves/hour (sharded sweep)", "n_gpus": world, "system": a.system, "nint": a.nint,
            "guesses": a.guesses, "chunk": a.chunk, "seconds": dt, "solves_per_hour": a.guesses / dt * 3600,
            "converged": int(conv.sum()), "iteration_limit": int((st == 2).sum()), "non_finite": int((st == 3).sum()),
            "distinct_actions_6_digits": int(np.unique(acts).size), "U1_evals": nU1, "U2_evals": nU2,
            "newton_iterations_mean": float(merged["nit"].mean()), "scaling": "strong (fixed number of guesses)",
            "gathered_records": int(merged["index"].size), "collective": "none on the data path; gather_object of the records"}))
    if world > 1:
        dist.barrier(); dist.destroy_process_group()


if __name__ == "__main__":
    main()
