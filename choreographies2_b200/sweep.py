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
    With ``world == 1`` (or no store) it degenerates to a local counter."""

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
