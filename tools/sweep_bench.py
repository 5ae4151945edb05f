#!/usr/bin/env python
"""Sharded orbit-search sweep (BASELINE config 5): N independent initial guesses over the ranks (one process per GPU, plan
replicated, NO data-path collective).  Every rank claims chunks of seeds from a dynamic queue (sweep.ChunkQueue: one atomic
counter in a host-side TCPStore -- solves have variable iteration counts), runs the reference's search loop on each chunk
(refine.find_batched: wolfe, six optimiser levels, refinement x2; settings of tests/NewSym_data/3C/choreo_config.json), hashes
the saved orbits on the device (segmpos_to_hash, all exponents in one launch) and filters duplicates inside the chunk; rank 0
gathers (seed, outcome, level, |grad|, action, hash) host-side, merges the duplicate groups across ranks (TestHashSame) and
prints one JSON line with solves/hour and the list of distinct orbits.

    python tools/sweep_bench.py --guesses 4096                                        # 1 GPU
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 tools/sweep_bench.py --guesses 32768
    ... --cpu-seeds 64     # also run the reference's loop on the first seeds (rank 0, all host cores) and compare the orbit lists
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def _cpu_worker(task):
    seeds, nint0, reconverge = task
    for k in ("OMP_NUM_THREADS", "MKL_NUM_THREADS", "OPENBLAS_NUM_THREADS"):
        os.environ[k] = "1"
    from oracle import ref_loader, ref_find
    ref_loader.import_reference()
    NBS = ref_loader.build_named("3C")
    fac0 = nint0 // (2 * int(NBS.nint_min))
    NBS.nint_fac = fac0
    x0 = ref_find.find_starts(NBS, seeds)
    cfg = {k: v for k, v in ref_find.CONFIG_3C.items() if k not in ("nint_fac_init", "coeff_ampl_o", "k_infl", "k_max", "coeff_ampl_min")}
    cfg["n_reconverge_it_max"] = reconverge
    out = []
    t = time.perf_counter()
    for i, s in enumerate(seeds):
        o, lvl, iopt, fn, ff, xb, nit = ref_find.reference_find(NBS, x0[i], fac0, **cfg)
        h = None
        if o in (1, 2):
            NBS.nint_fac = fac0 * 2 ** lvl
            h = np.asarray(NBS.segmpos_to_hash(NBS.params_to_segmpos(np.ascontiguousarray(xb)))).tolist()
        out.append((int(s), int(o), int(lvl), h))
    return time.perf_counter() - t, out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--system", default="3C")
    ap.add_argument("--nint", type=int, default=3072)
    ap.add_argument("--guesses", type=int, default=4096)
    ap.add_argument("--chunk", type=int, default=2048)
    ap.add_argument("--static", action="store_true", help="static contiguous split instead of the dynamic chunk queue (A/B)")
    ap.add_argument("--cpu-seeds", type=int, default=0, help="seeds also run through the reference's loop on the host (rank 0)")
    ap.add_argument("--reconverge", type=int, default=4, help="n_reconverge_it_max of both arms (the 3C configuration has 4)")
    ap.add_argument("--duplicate-eps", type=float, default=1e-8, help="rtol of TestHashSame (Solver_Checks.duplicate_eps)")
    ap.add_argument("--workers", type=int, default=1, help="chunks in flight per GPU (threads, one stream and one set of plans each)")
    ap.add_argument("--trace", default=None, help="write one line per solved (refinement level, optimiser level) group of rank 0 to this file")
    a = ap.parse_args()
    import torch
    import torch.distributed as dist
    from choreographies2_b200 import NBodySystB200, PlanTables, sweep, refine
    rank, world, local = sweep.env_rank_world()
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    T = PlanTables.load(os.path.join(ROOT, "choreographies2_b200", "systems", f"{a.system}.npz"))
    W = max(1, a.workers)
    kw = dict(refine.SEARCH_CONFIG_3C)
    kw["n_reconverge_it_max"] = a.reconverge
    # warm-up: plans of the levels, workspaces (a private queue: the shared counter is for the timed sweep)
    warm = sweep.run_search(T, a.nint, 128 * W, device=local, chunk=128, workers=W, config=kw, queue=sweep.ChunkQueue(128 * W, 128))
    plans = warm["plans"]

    def make(nint):  # plans of worker 0, for the merge on rank 0
        if nint not in plans[0]:
            plans[0][nint] = NBodySystB200(T, nint, device=local)
        return plans[0][nint]

    if a.static:
        lo_s, hi_s = sweep.shard_bounds(a.guesses, rank, world)
        queue, off, nmine = sweep.ChunkQueue(hi_s - lo_s, a.chunk), lo_s, hi_s - lo_s  # this rank's contiguous share, claimed locally
    else:
        queue, off, nmine = sweep.ChunkQueue(a.guesses, a.chunk, rank, world), 0, a.guesses
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    trace = [] if (a.trace and rank == 0) else None
    t0 = time.perf_counter()
    out_s = sweep.run_search(T, a.nint, nmine, device=local, chunk=a.chunk, workers=W, config=kw, queue=queue, seed_offset=off,
                             duplicate_eps=a.duplicate_eps, trace=trace, plans=plans)
    rec = out_s["records"]
    nU1, nU2, nchunks, t_starts = out_s["U1"], out_s["U2"], out_s["chunks"], out_s["starts_seconds"]
    torch.cuda.synchronize()
    dt_local = time.perf_counter() - t0
    if trace is not None:
        with open(a.trace, "w") as fh:
            for r in trace:
                fh.write(json.dumps(r) + "\n")
    dt = dt_local
    if world > 1:
        t = torch.tensor([dt_local, float(nU1), float(nU2), float(nchunks)], dtype=torch.float64, device=dev)
        tmax, tmin = t.clone(), t.clone()
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        dist.all_reduce(tmin, op=dist.ReduceOp.MIN)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        dt, nU1, nU2 = float(tmax[0]), int(t[1]), int(t[2])
        balance = {"seconds_min_rank": float(tmin[0]), "seconds_max_rank": float(tmax[0]), "chunks_min": int(tmin[3]), "chunks_max": int(tmax[3])}
    else:
        balance = {"seconds_min_rank": dt, "seconds_max_rank": dt, "chunks_min": nchunks, "chunks_max": nchunks}
    local_rec = {k: rec[k] for k in ("outcome", "level", "gradnorm", "action", "hash", "dup_of_seed")}
    merged = sweep.gather_records(local_rec, rec["seed"])
    if rank == 0:
        out = merged["outcome"]
        saved = np.nonzero((out == refine.FOUND) | (out == refine.APPROX))[0]
        # cross-chunk / cross-rank merge: the representatives of the chunks against each other
        reps = saved[merged["dup_of_seed"][saved] == merged["index"][saved]]
        t1 = time.perf_counter()
        grp = refine.group_duplicates(make(a.nint), merged["hash"][reps], a.duplicate_eps)
        t_merge = time.perf_counter() - t1
        distinct = reps[np.unique(grp)]
        orbits = [{"seed": int(merged["index"][i]), "action": float(merged["action"][i]), "hash0": float(merged["hash"][i, 0]),
                   "level": int(merged["level"][i])} for i in distinct]
        line = {
            "metric": "Newton orbit solves/hour (sharded search sweep with refinement, hash and duplicate filter)", "n_gpus": world,
            "system": a.system, "nint0": a.nint, "n_reconverge_it_max": a.reconverge, "guesses": a.guesses, "chunk": a.chunk, "workers_per_gpu": W, "queue": "static split" if a.static else "dynamic (TCPStore counter)",
            "seconds": dt, "solves_per_hour": a.guesses / dt * 3600, "balance": balance,
            "found": int((out == refine.FOUND).sum()), "approximate": int((out == refine.APPROX).sum()),
            "not_converged": int((out == refine.NOT_CONVERGED).sum()), "collision": int((out == refine.COLLISION).sum()),
            "escaped": int((out == refine.ESCAPED).sum()), "refined": int((merged["level"] > 0).sum()),
            "chunk_representatives": int(reps.size), "distinct_orbits": int(distinct.size), "merge_seconds_rank0": t_merge,
            "duplicates": int(saved.size - distinct.size), "U1_evals": nU1, "U2_evals": nU2, "starts_seconds_rank0": t_starts,
            "scaling": "strong (fixed number of guesses)", "collective": "none on the data path; TCPStore counter + gather_object of the records",
            "orbits": sorted(orbits, key=lambda o: o["action"])[:64],
        }
        if a.cpu_seeds > 0:
            import multiprocessing as mp
            P = min(os.cpu_count() or 1, a.cpu_seeds)
            per = (a.cpu_seeds + P - 1) // P
            tasks = [(list(range(p * per, min(a.cpu_seeds, (p + 1) * per))), a.nint, a.reconverge) for p in range(P)]
            tasks = [t for t in tasks if t[0]]
            tc = time.perf_counter()
            with mp.get_context("spawn").Pool(len(tasks)) as pool:
                resc = pool.map(_cpu_worker, tasks, chunksize=1)
            wall = time.perf_counter() - tc
            busy = max(r[0] for r in resc)
            rows = sorted(r for res_ in resc for r in res_[1])
            same_outcome = sum(1 for (s, o, lvl, h) in rows if int(out[s]) == o and int(merged["level"][s]) == lvl)
            nbs0 = make(a.nint)
            same_orbit = n_cpu_saved = 0
            for (s, o, lvl, h) in rows:
                if o in (1, 2):
                    n_cpu_saved += 1
                    if int(out[s]) == o and nbs0.TestHashSame(np.asarray(h), merged["hash"][s], a.duplicate_eps):
                        same_orbit += 1
            cpu_h = np.array([h for (_, o, _, h) in rows if o in (1, 2)])
            cpu_distinct = int(np.unique(refine.group_duplicates(nbs0, cpu_h, a.duplicate_eps)).size) if len(cpu_h) else 0
            gsel = [s for (s, o, lvl, h) in rows if int(out[s]) in (1, 2)]
            gpu_distinct = int(np.unique(refine.group_duplicates(nbs0, merged["hash"][gsel], a.duplicate_eps)).size) if gsel else 0
            line["cpu_reference"] = {"seeds": len(rows), "processes": len(tasks), "seconds_busiest_process": busy, "seconds_wall": wall,
                                     "solves_per_hour": len(rows) / busy * 3600, "saved": n_cpu_saved, "same_outcome_and_level": same_outcome,
                                     "same_orbit_by_hash": same_orbit, "distinct_orbits_cpu": cpu_distinct, "distinct_orbits_b200_same_seeds": gpu_distinct}
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
