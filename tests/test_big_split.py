"""Residue-split staged transforms (choreographies2_b200/csrc/chb_big.cuh): every split R = 2, 4, 8 of the time index gives
the oracle's segment positions, action, gradient and Hessian-vector product; the splits agree with each other; the split picked
per launch does not change results.  Both backends (g++ emulation of the CTA programs on CPU, the CUDA library with -m gpu)."""
import os

import numpy as np
import pytest

from conftest import ROOT, TOL_ACTION, TOL_GRAD, TOL_HESS, TOL_TRANSFORM, make_nbs, relerr
from choreographies2_b200 import PlanTables
from oracle.choreo_oracle import OracleSyst


def _tables(name):
    return PlanTables.load(os.path.join(ROOT, "choreographies2_b200", "systems", f"{name}.npz"))


# (system, nint): one loop of three bodies (3C), three loops (3B), twenty loops with different lengths (20B), a 3 x 2^k length
CASES = [("3C", 768), ("3C", 3 * 2 * 80), ("3B", 256), ("20B", 240), ("16C", 512)]


@pytest.mark.parametrize("name,nint", CASES)
def test_every_split_matches_the_oracle(name, nint, backend):
    T = _tables(name)
    nbs = make_nbs(T, nint, backend)
    nbs.set_fused(False)
    avail = nbs.staged_kernels[0]
    assert avail & 2, "the level must have the parity split"
    O = OracleSyst(T, nint, backend="fft")
    rng = np.random.default_rng(7)
    x, dx = rng.random((3, nbs.nparams)), rng.random((3, nbs.nparams))
    seg_o = np.stack([O.params_to_segmpos(x[i]) for i in range(3)])
    act_o = np.array([O.params_to_action(x[i]) for i in range(3)])
    grad_o = np.stack([O.params_to_action_grad(x[i]) for i in range(3)])
    hess_o = np.stack([O.params_to_action_hess(x[i], dx[i]) for i in range(3)])
    parT_o = np.stack([O.segmpos_to_params_T(seg_o[i]) for i in range(3)])
    seen = {}
    for R in (2, 4, 8):
        if not avail & R:
            continue
        nbs.set_transform_split(R)
        tag = f"{name}/{nint}/R={R}"
        seg = nbs.params_to_segmpos(x)
        act, grad = nbs.params_to_action_and_grad(x)
        hess = nbs.params_to_action_hess(x, dx)
        parT = nbs.segmpos_to_params_T(seg_o)
        assert relerr(seg, seg_o) < TOL_TRANSFORM, tag
        assert relerr(act, act_o) < TOL_ACTION, tag
        assert relerr(grad, grad_o) < TOL_GRAD, tag
        assert relerr(hess, hess_o) < TOL_HESS, tag
        assert relerr(parT, parT_o) < TOL_TRANSFORM, tag
        assert relerr(nbs.params_to_kin_nrg(x), [O.params_to_kin_nrg(x[i]) for i in range(3)]) < TOL_ACTION, tag
        seen[R] = (seg, grad)
    assert len(seen) >= 2, f"{name}/{nint}: expected more than one split (available {avail})"
    base = seen[2]
    for R, (seg, grad) in seen.items():
        assert relerr(seg, base[0]) < 1e-13 and relerr(grad, base[1]) < 1e-12, R
    nbs.set_transform_split(0)  # automatic: a batch of 3 states takes the widest split, same numbers
    act, grad = nbs.params_to_action_and_grad(x)
    assert relerr(act, act_o) < TOL_ACTION and relerr(grad, grad_o) < TOL_GRAD


def test_split_argument_is_validated(backend):
    nbs = make_nbs(_tables("3C"), 768, backend)
    with pytest.raises(Exception):
        nbs.set_transform_split(3)
    nbs.set_transform_split(0)
