"""Multi-process (gloo, world_size 2 and 3) tests of the batch sharding + host-side gather.

The evaluator on each rank is the real host layer (NBodySystB200) driven over the g++ host emulation of the CTA programs
(tests/hostsim) -- the build container has no GPU -- so the N>1 path is exercised end to end: shard bounds, per-rank
plan, per-rank evaluation, gather, and the result must equal the single-process evaluation bit for bit."""
import os
import socket
import sys

import numpy as np
import pytest

from conftest import ROOT
from choreographies2_b200 import sweep


def test_shard_bounds_cover_and_balance():
    for n in (0, 1, 7, 8, 4096, 32768 + 5):
        for world in (1, 2, 3, 8):
            b = [sweep.shard_bounds(n, r, world) for r in range(world)]
            assert b[0][0] == 0 and b[-1][1] == n
            assert all(b[i][1] == b[i + 1][0] for i in range(world - 1))
            sizes = [hi - lo for lo, hi in b]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        sweep.shard_bounds(4, 2, 2)


def test_chunk_schedule_round_robin():
    s = sweep.chunk_schedule(10, 3, 4)
    assert s == [(0, 0, 4), (1, 4, 8), (2, 8, 10)]
    owners = [o for o, _, _ in sweep.chunk_schedule(32768, 8, 512)]
    assert all(owners.count(r) == 8 for r in range(8))


def test_gather_single_process():
    out = sweep.gather_records({"a": np.array([3.0, 1.0])}, np.array([5, 2]))
    assert out["index"].tolist() == [2, 5] and out["a"].tolist() == [1.0, 3.0]
    with pytest.raises(ValueError):
        sweep.gather_records({"a": np.zeros(3)}, np.array([0, 1]))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n, name, nint, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch.distributed as dist
    from choreographies2_b200 import NBodySystB200, PlanTables, sweep as sw
    from hostsim import build_hostsim
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        T = PlanTables.load(os.path.join(ROOT, "choreographies2_b200", "systems", f"{name}.npz"))
        nbs = NBodySystB200(T, nint, _lib_handle=build_hostsim.load())

        def maker(lo, hi):  # seeded by global state index: independent of the number of ranks
            return np.stack([np.random.default_rng(1000 + i).random(nbs.nparams) for i in range(lo, hi)]) \
                if hi > lo else np.zeros((0, nbs.nparams))

        def evaluate(x):
            act, grad = nbs.params_to_action_and_grad(x)
            return {"action": act, "gradnorm": np.linalg.norm(grad, axis=1), "grad0": grad[:, 0]}

        r, w, _ = sw.env_rank_world()
        out = sw.sharded_eval(evaluate, maker, n, r, w)
        q.put((rank, out))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n", [(2, 5), (3, 2)])
def test_sharded_eval_gloo_matches_single_process(world, n):
    import multiprocessing as mp
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from choreographies2_b200 import NBodySystB200, PlanTables
    from hostsim import build_hostsim
    name, nint = "3C", 3 * 2 * 16
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n, name, nint, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=180) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(got[r] is None for r in range(1, world))
    out = got[0]
    assert out["index"].tolist() == list(range(n))
    # single-process evaluation of the same seeded states
    T = PlanTables.load(os.path.join(ROOT, "choreographies2_b200", "systems", f"{name}.npz"))
    nbs = NBodySystB200(T, nint, _lib_handle=build_hostsim.load())
    x = np.stack([np.random.default_rng(1000 + i).random(nbs.nparams) for i in range(n)])
    act, grad = nbs.params_to_action_and_grad(x)
    assert np.array_equal(out["action"], act)
    assert np.array_equal(out["grad0"], grad[:, 0])
    assert np.array_equal(out["gradnorm"], np.linalg.norm(grad, axis=1))


def _queue_worker(rank, world, port, n, chunk, delay, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import time
    import torch.distributed as dist
    from choreographies2_b200 import NBodySystB200, PlanTables, sweep as sw, refine
    from hostsim import build_hostsim
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        T = PlanTables.load(os.path.join(ROOT, "choreographies2_b200", "systems", "3C.npz"))
        lib = build_hostsim.load()
        nint = 3 * 2 * 16
        queue = sw.ChunkQueue(n, chunk, rank, world)
        mine, outcome = [], []
        while True:
            c = queue.next()
            if c is None:
                break
            time.sleep(delay * rank)  # rank 1 is the slow one: the fast rank must end up with more chunks
            lo, hi = c
            x0 = refine.find_starts(T, nint, range(lo, hi))
            res = refine.find_batched(lambda ni: NBodySystB200(T, ni, _lib_handle=lib), x0, nint, gradtol=1e-9, maxiter=30,
                                      n_reconverge_it_max=0, Newt_err_norm_max=1e-7, Newt_err_norm_max_save=1e-5)
            mine.append((lo, hi))
            outcome.append(res.outcome)
        idx = np.concatenate([np.arange(lo, hi) for lo, hi in mine]) if mine else np.zeros(0, dtype=np.int64)
        rec = {"outcome": np.concatenate(outcome) if outcome else np.zeros(0, dtype=np.int64)}
        q.put((rank, mine, sw.gather_records(rec, idx)))
    finally:
        dist.destroy_process_group()


def test_dynamic_chunk_queue_gloo():
    """sweep.ChunkQueue over two gloo ranks: every chunk is claimed exactly once, the slower rank takes fewer chunks, and the
    gathered per-seed outcomes equal a single-process run (the result does not depend on who solved which chunk)."""
    import multiprocessing as mp
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from choreographies2_b200 import NBodySystB200, PlanTables, refine
    from hostsim import build_hostsim
    n, chunk, world = 14, 2, 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_queue_worker, args=(r, world, port, n, chunk, 0.4, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = {r: (mine, merged) for r, mine, merged in (q.get(timeout=300) for _ in range(world))}
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    claimed = sorted(got[0][0] + got[1][0])
    assert claimed == [(lo, min(n, lo + chunk)) for lo in range(0, n, chunk)]
    assert len(got[0][0]) > len(got[1][0]) >= 1
    merged = got[0][1]
    assert got[1][1] is None and merged["index"].tolist() == list(range(n))
    T = PlanTables.load(os.path.join(ROOT, "choreographies2_b200", "systems", "3C.npz"))
    lib = build_hostsim.load()
    nint = 3 * 2 * 16
    res = refine.find_batched(lambda ni: NBodySystB200(T, ni, _lib_handle=lib), refine.find_starts(T, nint, range(n)), nint, gradtol=1e-9,
                              maxiter=30, n_reconverge_it_max=0, Newt_err_norm_max=1e-7, Newt_err_norm_max_save=1e-5)
    assert np.array_equal(merged["outcome"], res.outcome)
    # single-process queue
    from choreographies2_b200 import sweep as sw
    q1 = sw.ChunkQueue(5, 2)
    assert [q1.next(), q1.next(), q1.next(), q1.next()] == [(0, 2), (2, 4), (4, 5), None]


def _search_worker(rank, world, port, n, chunk, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch.distributed as dist
    from choreographies2_b200 import NBodySystB200, PlanTables, sweep as sw, refine
    from hostsim import build_hostsim
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        T = PlanTables.load(os.path.join(ROOT, "choreographies2_b200", "systems", "3C.npz"))
        lib = build_hostsim.load()
        cfg = dict(refine.SEARCH_CONFIG_3C)
        cfg["n_reconverge_it_max"] = 1
        queue = sw.ChunkQueue(n, chunk, rank, world)  # rank 0 hosts the counter: the object must outlive the other ranks' claims
        out = sw.run_search(T, 96, n, device=None, chunk=chunk, workers=2, config=cfg, queue=queue,
                            make_plan=lambda ni: NBodySystB200(T, ni, _lib_handle=lib))
        rec = out["records"]
        merged = sw.gather_records({k: rec[k] for k in ("outcome", "level", "action", "dup_of_seed")}, rec["seed"])
        q.put((rank, int(out["chunks"]), rec["seed"].tolist(), merged))
        del queue  # after the gather (a collective): every rank has finished claiming
    finally:
        dist.destroy_process_group()


def test_run_search_two_ranks_gloo():
    """sweep.run_search on two gloo ranks sharing the TCPStore chunk queue, two worker threads per rank: every seed is solved
    exactly once, and the merged records equal a single-process run (no collective on the data path, gather at the end)."""
    import multiprocessing as mp
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from choreographies2_b200 import NBodySystB200, PlanTables, refine, sweep as sw
    from hostsim import build_hostsim
    n, chunk, world = 12, 2, 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_search_worker, args=(r, world, port, n, chunk, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = {r: (nch, seeds, merged) for r, nch, seeds, merged in (q.get(timeout=300) for _ in range(world))}
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert got[0][0] + got[1][0] == n // chunk
    assert sorted(got[0][1] + got[1][1]) == list(range(n))
    merged = got[0][2]
    assert got[1][2] is None and merged["index"].tolist() == list(range(n))
    T = PlanTables.load(os.path.join(ROOT, "choreographies2_b200", "systems", "3C.npz"))
    lib = build_hostsim.load()
    cfg = dict(refine.SEARCH_CONFIG_3C)
    cfg["n_reconverge_it_max"] = 1
    one = sw.run_search(T, 96, n, device=None, chunk=n, workers=1, config=cfg, make_plan=lambda ni: NBodySystB200(T, ni, _lib_handle=lib))["records"]
    assert np.array_equal(merged["outcome"], one["outcome"]) and np.array_equal(merged["level"], one["level"])
    saved = (one["outcome"] == refine.FOUND) | (one["outcome"] == refine.APPROX)
    assert np.allclose(merged["action"][saved], one["action"][saved], rtol=1e-12, atol=0)
