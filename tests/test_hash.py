"""segmpos_to_hash: the library's pair-sum kernel with the law (x^2)^e against a NumPy restatement of
segm_pos_to_hash (_NBodySyst.pyx:6271-6369) and, when oracle/_ref is built, against the unmodified reference."""
import os
import sys

import numpy as np
import pytest

from conftest import ROOT, golden_names, relerr
from choreographies2_b200 import NBodySystB200, PlanTables


def _hash_numpy(T, nint, segmpos, exps):
    lvl = T.level(nint)
    size, store, G = lvl.segm_size, lvl.segm_store, T.geodim
    w = np.ones(store)
    if store != size:
        w[0] = w[-1] = 0.5
    out = np.zeros(len(exps))
    rot = np.asarray(T.bin_spacerot).reshape(T.nbin, G, G)
    for b in range(T.nbin):
        d = segmpos[int(T.bin_source[b])] @ rot[b].T - segmpos[int(T.bin_target[b])]
        dx2 = (d * d).sum(-1)
        for k, e in enumerate(exps):
            out[k] += float(T.bin_prodchargesum[b]) / size * float((w * dx2 ** e).sum())
    return out


def _check(lib_handle, names):
    for name in names:
        T = PlanTables.load(os.path.join(ROOT, "choreographies2_b200", "systems", f"{name}.npz"))
        nint = 2 * T.nint_min * 8
        nbs = NBodySystB200(T, nint) if lib_handle is None else NBodySystB200(T, nint, _lib_handle=lib_handle)
        rng = np.random.default_rng(4)
        x = rng.random((3, nbs.nparams))
        seg = nbs.params_to_segmpos(x)
        H = nbs.segmpos_to_hash(seg)
        assert H.shape == (3, 9)
        for i in range(3):
            assert relerr(H[i], _hash_numpy(T, nint, seg[i], nbs.Hash_exp)) < 1e-12, (name, i)
        assert relerr(nbs.segmpos_to_hash(seg[1]), H[1]) == 0.0
        assert nbs.TestHashSame(H[0], H[0]) and not nbs.TestHashSame(H[0], H[1])


def test_hash_hostsim_matches_numpy():
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from hostsim import build_hostsim
    _check(build_hostsim.load(), ["3C", "3D", "5Dq", "20B"])


def test_hash_matches_reference():
    from oracle import ref_loader
    if not ref_loader.reference_available():
        pytest.skip("oracle/_ref not built")
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from hostsim import build_hostsim
    ref_loader.import_reference()
    from choreographies2_b200 import extract_tables
    for name in ("3C", "5Dq", "3D"):
        NBS = ref_loader.build_named(name)
        NBS.nint = 2 * NBS.nint_min * 8
        nbs = NBodySystB200(extract_tables(NBS), NBS.nint, _lib_handle=build_hostsim.load())
        x = np.random.default_rng(1).random(NBS.nparams)
        seg = NBS.params_to_segmpos(x)
        assert np.allclose(NBS.Hash_exp, nbs.Hash_exp)
        Hr, Hb = np.asarray(NBS.segmpos_to_hash(seg)), nbs.segmpos_to_hash(seg)
        assert relerr(Hb, Hr) < 1e-12, name
        # a rescaled copy of the orbit is "the same" when multiples are detected (TestHashSame, :2036-2067)
        seg2 = NBS.params_to_segmpos(1.7 * x)
        H2r, H2b = np.asarray(NBS.segmpos_to_hash(seg2)), nbs.segmpos_to_hash(seg2)
        assert NBS.TestHashSame(Hr, H2r) == nbs.TestHashSame(Hb, H2b)
        assert NBS.TestHashSame(Hr, H2r, 1e-5, False) == nbs.TestHashSame(Hb, H2b, 1e-5, False)


@pytest.mark.gpu
def test_hash_cuda_matches_numpy():
    import torch
    _check(None, ["3C", "3D", "5Dq", "20B", "3B"])
    T = PlanTables.load(os.path.join(ROOT, "choreographies2_b200", "systems", "3C.npz"))
    nbs = NBodySystB200(T, 3072)
    x = torch.rand((64, nbs.nparams), dtype=torch.float64, device="cuda")
    seg = nbs.params_to_segmpos(x)
    Hd = nbs.segmpos_to_hash(seg).cpu().numpy()
    Hh = nbs.segmpos_to_hash(seg.cpu().numpy())
    assert relerr(Hd, Hh) < 1e-13
