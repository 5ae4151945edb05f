"""The one-launch Arnoldi step of the batched GMRES (chb_arnoldi_step_dev, chb_krylov.cuh) against a NumPy restatement of
the same step (shifted operator, Gram-Schmidt twice where needed), on the g++ host emulation here and on the GPU."""
import os
import sys

import numpy as np
import pytest

from conftest import ROOT
from choreographies2_b200 import NBodySystB200, PlanTables

sys.path.insert(0, os.path.join(ROOT, "tests"))


def _numpy_step(V, j, w, skip):
    """Reference: orthogonalise w - v_j against v_0..v_j in extended precision-free plain NumPy (two full passes)."""
    A, n = w.shape
    h = np.zeros((A, V.shape[0]))
    vnew = np.zeros((A, n))
    for a in range(A):
        if skip is not None and skip[a]:
            continue
        u = w[a] - V[j, a]
        for _ in range(2):
            c = V[:j + 1, a] @ u
            u = u - c @ V[:j + 1, a]
            h[a, :j + 1] += c
        h[a, j] += 1.0
        h[a, j + 1] = np.linalg.norm(u)
        vnew[a] = u / h[a, j + 1]
    return h, vnew


def _case(A, n, j, m, seed, with_skip, noise=1e-2):
    rng = np.random.default_rng(seed)
    V = np.zeros((m + 1, A, n))
    for a in range(A):  # orthonormal v_0..v_j per state
        q, _ = np.linalg.qr(rng.standard_normal((n, j + 1)))
        V[:j + 1, a] = q.T
    # A v_j = v_j + small perturbation with components along the basis (what the action Hessian produces)
    # (noise 1e-4: what is left after the first Gram-Schmidt pass is < 0.1 of the vector -> the second pass runs)
    w = V[j] + noise * rng.standard_normal((A, n)) + 0.3 * V[0]
    skip = (rng.random(A) < 0.3) if with_skip else None
    if skip is not None:
        skip[0] = False  # at least one running state
    return V, w, skip


def _check(nbs, to_dev, from_dev, A, n, j, m, seed, with_skip, noise=1e-2):
    V, w, skip = _case(A, n, j, m, seed, with_skip, noise)
    h_ref, v_ref = _numpy_step(V, j, w, skip)
    Vd, wd, hd = to_dev(V), to_dev(w), to_dev(np.zeros((A, m + 1)))
    sd = to_dev(skip) if skip is not None else None
    assert nbs.arnoldi_step(Vd, j, wd, sd, hd)
    h, Vn = from_dev(hd), from_dev(Vd)
    assert np.abs(h - h_ref).max() < 1e-12
    assert np.abs(Vn[j + 1] - v_ref).max() < 1e-11
    assert np.array_equal(Vn[:j + 1], V[:j + 1])
    act = np.ones(A, bool) if skip is None else ~skip
    # the new vector is a unit vector orthogonal to the basis
    G = np.einsum("jan,an->ja", Vn[:j + 1, act], Vn[j + 1, act])
    assert np.abs(G).max() < 1e-13
    assert np.abs(np.linalg.norm(Vn[j + 1, act], axis=1) - 1).max() < 1e-13


@pytest.mark.parametrize("n,j,with_skip", [(37, 0, False), (300, 3, True), (2049, 19, True)])
def test_arnoldi_step_hostsim(n, j, with_skip):
    from hostsim import build_hostsim
    T = PlanTables.load(os.path.join(ROOT, "choreographies2_b200", "systems", "3C.npz"))
    nbs = NBodySystB200(T, 96, _lib_handle=build_hostsim.load())
    _check(nbs, lambda a: np.ascontiguousarray(a), lambda a: a, 5, n, j, 20, 11 + n, with_skip)
    _check(nbs, lambda a: np.ascontiguousarray(a), lambda a: a, 3, n, j, 20, 12 + n, with_skip, noise=1e-4)


def test_arnoldi_step_declines_host_arrays_on_cuda_library():
    """NumPy arrays are host pointers: a build that reports CHB_Q_HOST_POINTERS = 0 (the CUDA library) must decline them, the
    driver then runs its array-level Gram-Schmidt."""
    nbs = NBodySystB200.__new__(NBodySystB200)
    nbs._q = lambda what: 0
    assert nbs.arnoldi_step(np.zeros((3, 2, 4)), 0, np.zeros((2, 4)), None, np.zeros((2, 3))) is False
    assert nbs.gmres_solution(np.zeros((3, 2, 4)), None, None, None, 1.0, None) is False
    assert nbs.hess_masked(np.zeros((2, 3)), None, None, None) is False


@pytest.mark.gpu
@pytest.mark.parametrize("n,j,with_skip", [(37, 0, False), (2000, 5, True), (4100, 7, False), (6142, 19, True),
                                           (6142, 2, False), (1534, 2, False), (8192, 3, True), (8193, 2, False), (24570, 4, True)])
def test_arnoldi_step_gpu(n, j, with_skip):
    import torch
    T = PlanTables.load(os.path.join(ROOT, "choreographies2_b200", "systems", "3C.npz"))
    nbs = NBodySystB200(T, 96)
    dev = torch.device("cuda:0")
    _check(nbs, lambda a: torch.as_tensor(np.ascontiguousarray(a), device=dev), lambda t: t.cpu().numpy(),
           7, n, j, 20, 5 + n, with_skip)
    _check(nbs, lambda a: torch.as_tensor(np.ascontiguousarray(a), device=dev), lambda t: t.cpu().numpy(),
           150, n, j, 20, 6 + n, with_skip, noise=1e-4)  # second Gram-Schmidt pass, more states than SMs


@pytest.mark.gpu
@pytest.mark.parametrize("n,j,with_skip", [(25601, 3, True), (30001, 6, False), (98302, 5, True), (204801, 2, False)])
def test_arnoldi_step_gpu_long_rows(n, j, with_skip):
    """Rows longer than one CTA's shared memory: the candidate vector distributed over a cluster of 8 CTAs (inner products through
    distributed shared memory), and beyond the cluster's capacity the global-memory variant."""
    import torch
    T = PlanTables.load(os.path.join(ROOT, "choreographies2_b200", "systems", "3C.npz"))
    nbs = NBodySystB200(T, 96)
    dev = torch.device("cuda:0")
    to_dev, from_dev = (lambda a: torch.as_tensor(np.ascontiguousarray(a), device=dev)), (lambda t: t.cpu().numpy())
    _check(nbs, to_dev, from_dev, 5, n, j, 8, 5 + n, with_skip)
    _check(nbs, to_dev, from_dev, 3, n, j, 8, 6 + n, with_skip, noise=1e-4)  # second Gram-Schmidt pass


def _masked_case(nbs, B, seed, forced):
    rng = np.random.default_rng(seed)
    x, dx = rng.random((B, nbs.nparams)), rng.random((B, nbs.nparams))
    skip = rng.random(B) < 0.4
    for b, v in forced.items():
        skip[b] = v
    return x, dx, skip


@pytest.mark.parametrize("name,nint,B,ver", [("3C", 96, 9, 3), ("3B", 4096, 300, 4)])
def test_hess_masked_hostsim(name, nint, B, ver):
    """The masked Hessian-vector operator of the GMRES loop: flagged states are left out by the fused kernels (their rows of the
    output keep what was there), the other rows equal the plain operator bit for bit.  The version-4 case has more states than
    the emulated grid has CTAs (148): CTA 0 starts on a flagged state, CTA 5 owns only flagged states, CTA 7 a flagged state
    between two live ones."""
    from hostsim import build_hostsim
    T = PlanTables.load(os.path.join(ROOT, "choreographies2_b200", "systems", f"{name}.npz"))
    nbs = NBodySystB200(T, nint, _lib_handle=build_hostsim.load())
    assert nbs.fused_version == ver
    forced = {0: True, 1: False}
    if B > 296:
        forced.update({5: True, 153: True, 7: False, 155: True, 148: False, 296: False, 2: False, 150: False, 298: True})
    x, dx, skip = _masked_case(nbs, B, 3, forced)
    live = np.nonzero(~skip)[0]
    seg = np.zeros((B,) + nbs.segmpos_shape)
    seg[live] = nbs.params_to_segmpos(x[live])
    out = np.full((B, nbs.nparams), -7.25)
    assert nbs.hess_masked(seg, dx, skip, out)
    ref = nbs.segmpos_dparams_to_action_hess(seg[live], dx[live])
    assert np.array_equal(out[live], ref)
    assert np.all(out[skip] == -7.25)
    # the same states given as a list (any order), longer than the emulated grid in the version-4 case
    lst = np.zeros(B, dtype=np.int32)
    lst[:live.size] = np.random.default_rng(0).permutation(live)
    out2 = np.full((B, nbs.nparams), -7.25)
    assert nbs.hess_listed(seg, dx, lst, np.array([live.size], dtype=np.int32), out2)
    assert np.array_equal(out2, out)
    out2[:] = 1.5
    assert nbs.hess_listed(seg, dx, lst, np.array([0], dtype=np.int32), out2) and np.all(out2 == 1.5)


def test_gmres_solution_hostsim():
    from hostsim import build_hostsim
    T = PlanTables.load(os.path.join(ROOT, "choreographies2_b200", "systems", "3C.npz"))
    nbs = NBodySystB200(T, 96, _lib_handle=build_hostsim.load())
    rng = np.random.default_rng(2)
    A, n, m = 6, 130, 20
    V = rng.standard_normal((m + 1, A, n))
    R = np.triu(rng.standard_normal((A, m + 1, m))) + 3 * np.eye(m + 1, m)
    g = rng.standard_normal((A, m + 1))
    kdone = np.array([0, 1, 5, 20, 13, 2])
    R[4, 3, 3] = 0.0  # zero pivot: replaced by 1
    out = np.empty((A, n))
    assert nbs.gmres_solution(V, R, g, kdone, -1.0, out)
    for a in range(A):
        k = kdone[a]
        Rk = R[a, :k, :k].copy()
        for i in range(k):
            if Rk[i, i] == 0:
                Rk[i, i] = 1.0
        y = np.linalg.solve(Rk, g[a, :k]) if k else np.zeros(0)
        assert np.abs(out[a] + y @ V[:k, a]).max() < 1e-9 * max(1.0, np.abs(y).max(initial=0.0))


@pytest.mark.gpu
@pytest.mark.parametrize("name,nint,B,ver", [("3C", 3072, 700, 3), ("3B", 4096, 1000, 4)])
def test_hess_masked_gpu(name, nint, B, ver):
    import torch
    T = PlanTables.load(os.path.join(ROOT, "choreographies2_b200", "systems", f"{name}.npz"))
    nbs = NBodySystB200(T, nint)
    assert nbs.fused_version == ver
    x, dx, skip = _masked_case(nbs, B, 4, {0: True, 1: False, B - 1: True})
    dev = torch.device("cuda:0")
    xd, dxd, sk = torch.as_tensor(x, device=dev), torch.as_tensor(dx, device=dev), torch.as_tensor(skip, device=dev)
    seg = nbs.params_to_segmpos(xd)
    out = torch.full((B, nbs.nparams), -7.25, dtype=torch.float64, device=dev)
    assert nbs.hess_masked(seg, dxd, sk, out)
    ref = nbs.segmpos_dparams_to_action_hess(seg, dxd)
    torch.cuda.synchronize()
    assert torch.equal(out[~sk], ref[~sk])
    assert bool((out[sk] == -7.25).all())
    live = torch.nonzero(~sk).flatten()
    lst = torch.zeros(B, dtype=torch.int32, device=dev)
    lst[:live.numel()] = live[torch.randperm(live.numel(), device=dev)].to(torch.int32)
    out2 = torch.full((B, nbs.nparams), -7.25, dtype=torch.float64, device=dev)
    assert nbs.hess_listed(seg, dxd, lst, torch.tensor([live.numel()], dtype=torch.int32, device=dev), out2)
    torch.cuda.synchronize()
    assert torch.equal(out2, out)
    assert nbs.hess_listed(seg, dxd, lst, torch.zeros(1, dtype=torch.int32, device=dev), out2)
    torch.cuda.synchronize()
    assert torch.equal(out2, out)
    # all states flagged: nothing is written, nothing hangs
    out.fill_(1.5)
    assert nbs.hess_masked(seg, dxd, torch.ones(B, dtype=torch.bool, device=dev), out)
    torch.cuda.synchronize()
    assert bool((out == 1.5).all())


@pytest.mark.gpu
def test_gmres_solution_gpu():
    import torch
    T = PlanTables.load(os.path.join(ROOT, "choreographies2_b200", "systems", "3C.npz"))
    nbs = NBodySystB200(T, 96)
    rng = np.random.default_rng(2)
    A, n, m = 9, 6142, 20
    V = rng.standard_normal((m + 1, A, n))
    R = np.triu(rng.standard_normal((A, m + 1, m))) + 3 * np.eye(m + 1, m)
    g = rng.standard_normal((A, m + 1))
    kdone = np.array([0, 1, 5, 20, 13, 2, 7, 19, 3])
    dev = torch.device("cuda:0")
    out = torch.empty((A, n), dtype=torch.float64, device=dev)
    assert nbs.gmres_solution(*(torch.as_tensor(a, device=dev) for a in (V, R, g, kdone)), 1.0, out)
    got = out.cpu().numpy()
    for a in range(A):
        k = kdone[a]
        y = np.linalg.solve(R[a, :k, :k], g[a, :k]) if k else np.zeros(0)
        assert np.abs(got[a] - y @ V[:k, a]).max() < 1e-9 * max(1.0, np.abs(y).max(initial=0.0))


def _row_stats_case():
    rng = np.random.default_rng(9)
    a = rng.standard_normal((7, 1531)) * np.logspace(-3, 3, 7)[:, None]
    a[2, 17] = np.nan
    a[4, 1530] = np.inf
    a[5, :] = 0.0
    return a


def _check_row_stats(ss, mx, a):
    ok = [0, 1, 3, 5, 6]
    assert np.allclose(ss[ok], (a[ok] ** 2).sum(-1), rtol=1e-13, atol=0)
    assert np.array_equal(mx[ok], np.abs(a[ok]).max(-1))
    assert not np.isfinite(ss[2]) and not np.isfinite(ss[4])


def test_row_stats_hostsim():
    from hostsim import build_hostsim
    T = PlanTables.load(os.path.join(ROOT, "choreographies2_b200", "systems", "3C.npz"))
    nbs = NBodySystB200(T, 96, _lib_handle=build_hostsim.load())
    a = _row_stats_case()
    ss, mx = nbs.row_stats(a)
    _check_row_stats(ss, mx, a)


@pytest.mark.gpu
def test_row_stats_gpu():
    import torch
    T = PlanTables.load(os.path.join(ROOT, "choreographies2_b200", "systems", "3C.npz"))
    nbs = NBodySystB200(T, 96)
    a = _row_stats_case()
    ss, mx = nbs.row_stats(torch.from_numpy(a).cuda())
    _check_row_stats(ss.cpu().numpy(), mx.cpu().numpy(), a)
    # a row's result does not depend on the batch it is in
    ss1, mx1 = nbs.row_stats(torch.from_numpy(a[3:4]).cuda())
    assert ss1.item() == ss[3].item() and mx1.item() == mx[3].item()
