"""Multi-GPU sharding of a batch of independent states.

Every state (initial guess / Newton iterate) is independent and the symmetry tables are tiny, so the partition is a
contiguous split of the batch (or of the seed range) over the ranks: one process per GPU, plan replicated per device,
**no data-path collective**.  The only communication is the end-of-sweep gather of per-state records
``(index, status, action, |grad|, ...)`` -- host side, through ``torch.distributed`` (``gloo`` on CPU, ``nccl`` on the
GPU box) -- which mirrors how the reference scales (independent workers, ``choreo/find.py:836-850``) and how it collects
solutions (``find.py:905-935``).

torch.distributed is plumbing here; nothing in this module touches the numerics.
"""
from __future__ import annotations

import os
import threading
from typing import Callable, Dict, Optional, Sequence, Tuple

import numpy as np


def shard_bounds(n: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous balanced split of ``range(n)``: the first ``n % world`` ranks own one extra item."""
    if world < 1 or not (0 <= rank < world) or n < 0:
        raise ValueError(f"shard_bounds: bad arguments n={n} rank={rank} world={world}")
    q, r = divmod(n, world)
    lo = rank * q + min(rank, r)
    return lo, lo + q + (1 if rank < r else 0)


def chunk_schedule(n: int, world: int, chunk: int) -> Sequence[Tuple[int, int, int]]:
    """Static round-robin schedule of seed chunks ``(owner_rank, lo, hi)`` for solves with uneven iteration counts:
    chunk ``c`` goes to rank ``c % world`` so that slow regions of the seed range are spread over the GPUs."""
    if chunk < 1:
        raise ValueError("chunk_schedule: chunk must be >= 1")
    out = []
    for c, lo in enumerate(range(0, n, chunk)):
        out.append((c % world, lo, min(n, lo + chunk)))
    return out


class ChunkQueue:
    """Dynamic work queue of seed chunks shared by the ranks (solves have variable iteration counts, so a rank
    that finishes early takes the next chunk instead of idling behind a static split).  The queue is ONE atomic counter in a
    ``torch.distributed.TCPStore`` on the host network (``add`` is atomic): no device collective, nothing on the data path.
    With ``world == 1`` (or no store) it degenerates to a local counter.  Rank 0 hosts the store: keep its ``ChunkQueue`` alive
    until every rank has finished claiming (e.g. until after the end-of-sweep gather), or the others lose the counter."""

    def __init__(self, n: int, chunk: int, rank: int = 0, world: int = 1, host: str = "127.0.0.1", port: Optional[int] = None,
                 key: str = "chb_next_chunk"):
        if chunk < 1 or n < 0:
            raise ValueError("ChunkQueue: need n >= 0 and chunk >= 1")
        self.n, self.chunk, self.key = int(n), int(chunk), key
        self.nchunks = (self.n + self.chunk - 1) // self.chunk
        self._local = 0
        self._lock = threading.Lock()  # several worker threads of one rank may share the queue
        self.store = None
        if world > 1:
            from torch.distributed import TCPStore
            if port is None:
                port = int(os.environ.get("MASTER_PORT", "29500")) + 23
            self.store = TCPStore(host, int(port), world, is_master=(rank == 0), wait_for_workers=True)

    def next(self) -> Optional[Tuple[int, int]]:
        """Claim the next unclaimed chunk: ``(lo, hi)`` seed bounds, or None when the queue is empty."""
        with self._lock:
            if self.store is None:
                c = self._local
                self._local += 1
            else:
                c = int(self.store.add(self.key, 1)) - 1
        if c >= self.nchunks:
            return None
        lo = c * self.chunk
        return lo, min(self.n, lo + self.chunk)


def env_rank_world() -> Tuple[int, int, int]:
    """(rank, world, local_rank) from the torchrun environment (1 process when absent)."""
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")),
            int(os.environ.get("LOCAL_RANK", "0")))


def gather_records(local: Dict[str, np.ndarray], index: np.ndarray, group=None, dst: int = 0) -> Optional[Dict[str, np.ndarray]]:
    """Host-side end-of-sweep gather.  ``local[k]`` are arrays whose leading dimension matches ``index`` (the global
    state / seed numbers this rank owns).  Rank ``dst`` gets the records of all ranks merged and sorted by global index
    (plus ``"index"``); the other ranks get ``None``.  Works without an initialised process group (single process)."""
    import torch.distributed as dist

    index = np.asarray(index, dtype=np.int64)
    for k, v in local.items():
        if np.asarray(v).shape[:1] != index.shape:
            raise ValueError(f"gather_records: field {k!r} has leading shape {np.asarray(v).shape[:1]}, expected {index.shape}")
    payload = {"index": index, **{k: np.ascontiguousarray(v) for k, v in local.items()}}
    if not (dist.is_available() and dist.is_initialized()):
        parts = [payload]
        rank = dst
    else:
        rank = dist.get_rank(group)
        world = dist.get_world_size(group)
        parts = [None] * world if rank == dst else None
        dist.gather_object(payload, parts, dst=dst, group=group)
    if rank != dst:
        return None
    keys = list(payload.keys())
    merged = {k: np.concatenate([np.asarray(p[k]) for p in parts], axis=0) for k in keys}
    order = np.argsort(merged["index"], kind="stable")
    if np.unique(merged["index"]).size != merged["index"].size:
        raise ValueError("gather_records: a state index is owned by more than one rank")
    return {k: v[order] for k, v in merged.items()}


def sharded_eval(evaluate: Callable[[np.ndarray], Dict[str, np.ndarray]], x_all_or_maker, n: int, rank: int, world: int,
                 group=None, dst: int = 0) -> Optional[Dict[str, np.ndarray]]:
    """Evaluate ``n`` independent states split contiguously over ``world`` ranks and gather the per-state records.

    ``x_all_or_maker``: either the full ``[n, nparams]`` array (every rank slices its shard) or a callable
    ``maker(lo, hi) -> [hi-lo, nparams]`` that produces only this rank's shard (seeded by global index, so the result does
    not depend on ``world``).  ``evaluate(x_shard)`` returns a dict of arrays with leading dimension ``hi - lo``
    (e.g. ``{"action": ..., "gradnorm": ...}`` from ``NBodySystB200.params_to_action_and_grad``)."""
    lo, hi = shard_bounds(n, rank, world)
    x = x_all_or_maker(lo, hi) if callable(x_all_or_maker) else np.asarray(x_all_or_maker)[lo:hi]
    rec = evaluate(x) if hi > lo else {}
    if hi == lo:  # an empty shard still takes part in the gather with correctly shaped empties
        probe = evaluate(x_all_or_maker(0, 1) if callable(x_all_or_maker) else np.asarray(x_all_or_maker)[0:1]) if n > 0 else {}
        rec = {k: np.asarray(v)[:0] for k, v in probe.items()}
    return gather_records(rec, np.arange(lo, hi, dtype=np.int64), group=group, dst=dst)


def run_search(tables, nint0: int, n_guesses: int, *, device, chunk: int = 4096, workers: int = 1, config: Optional[dict] = None,
               random_config: Optional[dict] = None, queue: Optional[ChunkQueue] = None, seed_offset: int = 0,
               duplicate_eps: float = 1e-8, nhash: int = 9, make_plan: Optional[Callable] = None, trace: Optional[list] = None,
               plans: Optional[list] = None) -> Dict[str, object]:
    """This rank's share of a search sweep (BASELINE configuration 5): seeds ``seed_offset + [0, n_guesses)`` are claimed in chunks from
    ``queue`` (default: a private ``ChunkQueue`` over all of them), every chunk goes through the reference's search loop on the GPU
    (``refine.find_batched`` with ``config``, default the 3C ``Solver_*`` settings), the saved orbits are hashed on the device
    (one launch for all exponents) and duplicates inside the chunk are grouped (``Check_Duplicates``).

    ``workers`` > 1: that many chunks in flight on this GPU, one host thread, one CUDA stream and one set of plans each (a plan
    serves one stream at a time).  The searches of a chunk end in long tails of a few states (a start on a collision course runs
    thousands of Newton iterations at the deepest refinement level); with several chunks in flight the tail of one chunk runs under
    the bulk of the others.  Results do not depend on ``workers`` or ``chunk`` (states are independent) except for the order of
    the records, which are returned sorted by seed.  ``device=None``: NumPy arrays on the host (tests on the g++ emulation of the
    CTA programs; ``make_plan`` is then required).  ``plans``: the ``"plans"`` entry of an earlier call (warm plans, workspaces).

    Returns ``{"records": {seed, outcome, level, gradnorm, action, hash, dup_of_seed}, "U1", "U2", "chunks", "starts_seconds",
    "plans", "launches"}``; ``dup_of_seed[i]`` is the seed of the chunk representative of a saved orbit (itself when it is the
    first of its kind), -1 when the start was not saved.  No collective: merge the ranks' records with ``gather_records``."""
    import time
    from . import refine
    from .nbodysyst import NBodySystB200

    on_gpu = device is not None
    if on_gpu:
        import torch
        dev = torch.device("cuda", device) if isinstance(device, int) else device
    elif make_plan is None:
        raise ValueError("run_search: device=None needs make_plan")
    W = max(1, int(workers))
    kw = dict(refine.SEARCH_CONFIG_3C if config is None else config)
    rkw = dict(refine.RANDOM_CONFIG_3C if random_config is None else random_config)
    if queue is None:
        queue = ChunkQueue(n_guesses, chunk)
    if plans is None or len(plans) < W:
        plans = [dict() for _ in range(W)]

    def maker(w):
        def make(nint):
            if nint not in plans[w]:
                plans[w][nint] = make_plan(nint) if make_plan is not None else NBodySystB200(tables, nint, device=dev.index)
            return plans[w][nint]
        return make

    rec = {"seed": [], "outcome": [], "level": [], "gradnorm": [], "action": [], "hash": [], "dup_of_seed": []}
    count = {"U1": 0, "U2": 0, "chunks": 0, "starts_s": 0.0}
    lock = threading.Lock()
    host = (lambda t: t.cpu().numpy()) if on_gpu else np.asarray

    def work(w):
        mk = maker(w)
        while True:
            c = queue.next()
            if c is None:
                return
            lo, hi = c
            ts = time.perf_counter()
            x0 = refine.find_starts(tables, nint0, range(seed_offset + lo, seed_offset + hi), **rkw)
            if on_gpu:
                x0 = torch.from_numpy(x0).to(dev)
            t_st = time.perf_counter() - ts
            res = refine.find_batched(mk, x0, nint0, trace=trace if w == 0 else None, **kw)
            n = hi - lo
            H = np.full((n, nhash), np.nan)
            act = np.full(n, np.nan)
            saved = np.nonzero((res.outcome == refine.FOUND) | (res.outcome == refine.APPROX))[0]
            for lvl in np.unique(res.level[saved]):  # one plan per refinement level
                ids = saved[res.level[saved] == lvl]
                nbs = mk(nint0 * 2 ** int(lvl))
                sel = [res.params[int(i)] for i in ids]
                xs = torch.stack(sel) if on_gpu else np.stack(sel)
                seg = nbs.params_to_segmpos(xs)
                H[ids] = host(nbs.segmpos_to_hash(seg))
                act[ids] = host(nbs.params_to_action(xs))
            dup = np.full(n, -1, dtype=np.int64)
            if saved.size:
                rep = refine.group_duplicates(mk(nint0), H[saved], duplicate_eps)
                dup[saved] = seed_offset + lo + saved[rep]
            with lock:
                count["U1"] += res.n_grad_evals
                count["U2"] += res.n_hess_evals
                count["chunks"] += 1
                count["starts_s"] += t_st
                rec["seed"].append(np.arange(seed_offset + lo, seed_offset + hi))
                rec["outcome"].append(res.outcome)
                rec["level"].append(res.level)
                rec["gradnorm"].append(res.f_norm)
                rec["action"].append(act)
                rec["hash"].append(H)
                rec["dup_of_seed"].append(dup)

    if W == 1:
        work(0)
    else:
        errors = []

        def run(w):
            try:
                if on_gpu:
                    torch.cuda.set_device(dev)
                    with torch.cuda.stream(torch.cuda.Stream(device=dev)):
                        work(w)
                        torch.cuda.current_stream().synchronize()
                else:
                    work(w)
            except BaseException as e:  # noqa: BLE001 -- re-raised in the caller's thread
                errors.append(e)

        threads = [threading.Thread(target=run, args=(w,)) for w in range(W)]
        for th in threads:
            th.start()
        for th in threads:
            th.join()
        if errors:
            raise errors[0]
    widths = {"seed": (), "outcome": (), "level": (), "gradnorm": (), "action": (), "hash": (nhash,), "dup_of_seed": ()}
    out = {k: (np.concatenate(v) if v else np.zeros((0,) + widths[k])) for k, v in rec.items()}
    order = np.argsort(out["seed"], kind="stable")
    out = {k: v[order] for k, v in out.items()}
    out["seed"] = out["seed"].astype(np.int64)
    launches = sum(int(pl.kernel_launches) for d in plans for pl in d.values() if hasattr(pl, "kernel_launches"))
    return {"records": out, "U1": count["U1"], "U2": count["U2"], "chunks": count["chunks"], "starts_seconds": count["starts_s"],
            "plans": plans, "launches": launches}
