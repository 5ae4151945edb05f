"""ODE right-hand sides: Compute_forces(_vectorized) / Compute_grad_forces(_vectorized) of the library
(chb_ode.cuh) against a NumPy restatement of _NBodySyst.pyx:7531-7638 / 7641-7790 and, when oracle/_ref is built, against the
unmodified reference called once per vector."""
import os
import sys

import numpy as np
import pytest

from conftest import ROOT, relerr
from choreographies2_b200 import NBodySystB200, PlanTables

TOL = 1e-12


def _law(T, dx2):
    """(V', V'') of the interaction law at dx2 (gravity_pot :111-118, power_law_pot :125-133)."""
    if int(T.inter_law) == 0:
        a = dx2 ** -2.5
        return 0.5 * dx2 * a, -0.75 * a
    mnu, mnnm1, num2, alpha = (float(v) for v in T.inter_law_params)
    a = alpha * dx2 ** num2
    return mnu * dx2 * a, mnnm1 * a


def _forces_numpy(T, pos, dpos=None):
    G = T.geodim
    rot = np.asarray(T.bin_spacerot).reshape(T.nbin, G, G)
    x = pos.reshape(T.nsegm, G)
    F = np.zeros_like(x)
    if dpos is not None:
        d = dpos.reshape(T.nsegm, G, -1)
        gF = np.zeros_like(d)
    for b in range(T.nbin):
        s, t, R = int(T.bin_source[b]), int(T.bin_target[b]), rot[b]
        dx = R @ x[s] - x[t]
        v1, v2 = _law(T, float(dx @ dx))
        fac = -4.0 * float(T.bin_prodchargesum[b])
        df = fac * v1 * dx
        F[s] += R.T @ df
        F[t] -= df
        if dpos is not None:
            ddx = R @ d[s] - d[t]  # [G, ndof]
            ddf = 2.0 * fac * v2 * np.outer(dx, dx @ ddx) + fac * v1 * ddx
            gF[s] += R.T @ ddf
            gF[t] -= ddf
    return (F.reshape(-1), gF.reshape(dpos.shape)) if dpos is not None else F.reshape(-1)


def _make(name, lib_handle):
    T = PlanTables.load(os.path.join(ROOT, "choreographies2_b200", "systems", f"{name}.npz"))
    nint = 2 * T.nint_min * 4
    return T, (NBodySystB200(T, nint) if lib_handle is None else NBodySystB200(T, nint, _lib_handle=lib_handle))


def _check(lib_handle, names, nvec=5, ndofs=(1, 3)):
    for name in names:
        T, nbs = _make(name, lib_handle)
        vs = T.nsegm * T.geodim
        rng = np.random.default_rng(7)
        pos = rng.standard_normal((nvec, vs))
        F = nbs.Compute_forces_vectorized(np.zeros(nvec), pos)
        assert F.shape == (nvec, vs)
        for i in range(nvec):
            assert relerr(F[i], _forces_numpy(T, pos[i])) < TOL, (name, i)
        assert relerr(nbs.Compute_forces(0.0, pos[2]), F[2]) == 0.0
        for ndof in tuple(ndofs) + (2 * vs,):
            dpos = rng.standard_normal((nvec, vs, ndof))
            gF = nbs.Compute_grad_forces_vectorized(np.zeros(nvec), pos, dpos)
            assert gF.shape == dpos.shape
            for i in range(nvec):
                assert relerr(gF[i], _forces_numpy(T, pos[i], dpos[i])[1]) < TOL, (name, ndof, i)
            assert relerr(nbs.Compute_grad_forces(0.0, pos[1], dpos[1]), gF[1]) == 0.0
        # tangent forces are the directional derivative of the forces
        h = 1e-6 * rng.standard_normal(vs)
        lin = nbs.Compute_grad_forces(0.0, pos[0], h[:, None])[:, 0]
        fd = 0.5 * (nbs.Compute_forces(0.0, pos[0] + h) - nbs.Compute_forces(0.0, pos[0] - h))
        assert relerr(lin, fd) < 1e-6, name
        # empty batches
        assert nbs.Compute_forces_vectorized(np.zeros(0), np.zeros((0, vs))).shape == (0, vs)
        assert nbs.Compute_grad_forces_vectorized(np.zeros(0), np.zeros((0, vs)), np.zeros((0, vs, 2))).shape == (0, vs, 2)
        with pytest.raises(ValueError):
            nbs.Compute_forces(0.0, np.zeros(vs + 1))
        with pytest.raises(ValueError):
            nbs.Compute_grad_forces_vectorized(np.zeros(2), np.zeros((2, vs)), np.zeros((3, vs, 2)))


def _sim():
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from hostsim import build_hostsim
    return build_hostsim.load()


def test_forces_hostsim_matches_numpy():
    _check(_sim(), ["3C", "3D", "5Dq", "20B", "3B"])


@pytest.mark.parametrize("name", ["3C", "5Dq", "3D", "3B", "5q", "64C", "2D1_3dim", "20B"])
def test_forces_match_reference(name, backend):
    """Compute_forces / Compute_grad_forces per vector against the unmodified reference; runs on the emulation (CPU suite) and,
    with -m gpu, on the CUDA library (dense gather kernels for 20B / 64C, rotated interactions for 5q)."""
    from conftest import make_nbs, need_reference
    ref_loader = need_reference()
    from choreographies2_b200 import extract_tables
    if name == "64C" and backend is not None:
        pytest.skip("64C on the emulation takes minutes; covered on the GPU")
    NBS = ref_loader.build_named(name)
    NBS.nint = 2 * NBS.nint_min * 4
    nbs = make_nbs(extract_tables(NBS), NBS.nint, backend)
    vs = NBS.nsegm * NBS.geodim
    rng = np.random.default_rng(3)
    nvec = 4 if backend is not None else 40
    pos = rng.standard_normal((nvec, vs))
    dpos = rng.standard_normal((nvec, vs, min(2 * vs, 24)))
    F = nbs.Compute_forces_vectorized(np.zeros(nvec), pos)
    gF = nbs.Compute_grad_forces_vectorized(np.zeros(nvec), pos, dpos)
    # the reference's own Compute_grad_forces segfaults on interactions with a space rotation (5q; probed in this container),
    # so the tangent forces of such systems are checked against the finite-difference property in _check instead
    ref_grad_ok = bool(np.asarray(NBS.BinSpaceRotIsId).all())
    for i in range(0, nvec, max(1, nvec // 4)):
        assert relerr(F[i], np.asarray(NBS.Compute_forces(0.0, pos[i].copy()))) < TOL, (name, i)
        if ref_grad_ok:
            assert relerr(gF[i], np.asarray(NBS.Compute_grad_forces(0.0, pos[i].copy(), dpos[i].copy()))) < TOL, (name, i)


@pytest.mark.gpu
def test_forces_cuda_matches_numpy():
    import torch
    _check(None, ["3C", "3D", "5Dq", "20B", "3B", "64C"], nvec=300)
    for name in ("3C", "64C", "5Dq"):
        T, nbs = _make(name, None)
        vs = T.nsegm * T.geodim
        g = torch.Generator(device="cuda").manual_seed(5)
        pos = torch.randn((1000, vs), dtype=torch.float64, device="cuda", generator=g)
        dpos = torch.randn((1000, vs, 7), dtype=torch.float64, device="cuda", generator=g)
        l0 = nbs.kernel_launches
        Fd = nbs.Compute_forces_vectorized(None, pos)
        gFd = nbs.Compute_grad_forces_vectorized(None, pos, dpos)
        assert nbs.kernel_launches == l0 + 2
        Fh = nbs.Compute_forces_vectorized(None, pos.cpu().numpy())
        gFh = nbs.Compute_grad_forces_vectorized(None, pos.cpu().numpy(), dpos.cpu().numpy())
        assert relerr(Fd.cpu().numpy(), Fh) == 0.0 and relerr(gFd.cpu().numpy(), gFh) == 0.0
        for i in (0, 499, 999):
            Fi, gFi = _forces_numpy(T, Fh[i] * 0 + pos[i].cpu().numpy(), dpos[i].cpu().numpy())
            assert relerr(Fh[i], Fi) < TOL and relerr(gFh[i], gFi) < TOL, (name, i)


@pytest.mark.gpu
def test_forces_cuda_global_memory_variant():
    """CHB_ODE_NOSMEM=1: the kernel variant for vectors too long for a shared-memory tile (read once at library load)."""
    import subprocess
    code = ("import sys; sys.path.insert(0, %r); sys.path.insert(0, %r); import test_ode_forces as t; "
            "t._check(None, ['3C', '5Dq', '20B'], nvec=200)") % (ROOT, os.path.join(ROOT, "tests"))
    env = dict(os.environ, CHB_ODE_NOSMEM="1")
    r = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
