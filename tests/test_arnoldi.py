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


def _case(A, n, j, m, seed, with_skip):
    rng = np.random.default_rng(seed)
    V = np.zeros((m + 1, A, n))
    for a in range(A):  # orthonormal v_0..v_j per state
        q, _ = np.linalg.qr(rng.standard_normal((n, j + 1)))
        V[:j + 1, a] = q.T
    # A v_j = v_j + small perturbation with components along the basis (what the action Hessian produces)
    w = V[j] + 1e-2 * rng.standard_normal((A, n)) + 0.3 * V[0]
    skip = (rng.random(A) < 0.3) if with_skip else None
    return V, w, skip


def _check(nbs, to_dev, from_dev, A, n, j, m, seed, with_skip):
    V, w, skip = _case(A, n, j, m, seed, with_skip)
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


def test_arnoldi_step_declines_host_arrays_on_cuda_library():
    from choreographies2_b200 import _lib
    if not os.path.isfile(_lib.LIB_PATH):
        pytest.skip("CUDA library not built")

    class _Fake:  # only the attribute arnoldi_step looks at
        _name = _lib.LIB_PATH
    nbs = NBodySystB200.__new__(NBodySystB200)
    nbs._lib = _Fake()
    nbs._plan = None
    assert nbs.arnoldi_step(np.zeros((3, 2, 4)), 0, np.zeros((2, 4)), None, np.zeros((2, 3))) is False


@pytest.mark.gpu
@pytest.mark.parametrize("n,j,with_skip", [(37, 0, False), (2000, 5, True), (4100, 7, False), (6142, 19, True),
                                           (8192, 3, True), (8193, 2, False), (24570, 4, True)])
def test_arnoldi_step_gpu(n, j, with_skip):
    import torch
    T = PlanTables.load(os.path.join(ROOT, "choreographies2_b200", "systems", "3C.npz"))
    nbs = NBodySystB200(T, 96)
    dev = torch.device("cuda:0")
    _check(nbs, lambda a: torch.as_tensor(np.ascontiguousarray(a), device=dev), lambda t: t.cpu().numpy(),
           7, n, j, 20, 5 + n, with_skip)
