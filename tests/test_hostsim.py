"""CPU tests of the host logic and of the CTA programs' index arithmetic.

The build container has no GPU, so the kernels' CTA programs (choreographies2_b200/csrc/chb_core.cuh) are compiled
for the host by tests/hostsim (g++ -DCHB_HOSTSIM: a kernel launch becomes a sequential loop over blocks/threads) and
driven through the SAME C ABI and the SAME Python host layer as the CUDA library.  This is a pre-flight check of the
plan logic; the parity tests proper are the `-m gpu` tests (tests/test_gpu_parity.py)."""
import numpy as np
import pytest

from conftest import TOL_ACTION, TOL_GRAD, TOL_HESS, TOL_TRANSFORM, golden_names, relerr
from choreographies2_b200 import NBodySystB200, PlanTables
from hostsim import build_hostsim

NAMES = golden_names()


@pytest.fixture(scope="module")
def sim():
    return build_hostsim.load()


@pytest.mark.parametrize("name", NAMES)
def test_hostsim_matches_reference_golden(name, sim, golden_cache):
    g = golden_cache(name)
    for variant, nint in g.cases():
        nbs = NBodySystB200(PlanTables.from_dict(g.tables_dict(variant)), nint, _lib_handle=sim)
        v = lambda k: g.vec(variant, nint, k)  # noqa: E731
        tag = f"{name}/{variant}/{nint}"
        assert nbs.nparams == v("params").shape[0], tag
        assert nbs.segmpos_shape == v("segmpos").shape, tag
        assert relerr(nbs.params_to_segmpos(v("params")), v("segmpos")) < TOL_TRANSFORM, tag
        assert relerr(nbs.params_to_kin_nrg(v("params")), v("kin")) < TOL_ACTION, tag
        assert relerr(nbs.params_to_pot_nrg(v("params")), v("pot")) < TOL_ACTION, tag
        assert relerr(nbs.params_to_action(v("params")), v("action")) < TOL_ACTION, tag
        act, grad = nbs.params_to_action_and_grad(v("params"))
        assert relerr(act, v("action")) < TOL_ACTION, tag
        assert relerr(grad, v("grad")) < TOL_GRAD, tag
        if nbs.fused:  # everything above and below ran on the fused CTA program: repeat the chain on the staged one
            nbs.set_fused(False)
            assert relerr(nbs.params_to_segmpos(v("params")), v("segmpos")) < TOL_TRANSFORM, tag
            act, grad = nbs.params_to_action_and_grad(v("params"))
            assert relerr(act, v("action")) < TOL_ACTION and relerr(grad, v("grad")) < TOL_GRAD, tag
            assert relerr(nbs.params_to_action_hess(v("params"), v("dparams")), v("hess")) < TOL_HESS, tag
            assert relerr(nbs.segmpos_dparams_to_action_hess(v("segmpos"), v("dparams")), v("hess_cached")) < TOL_HESS, tag
            nbs.set_fused(True)
        assert relerr(nbs.params_to_action_hess(v("params"), v("dparams")), v("hess")) < TOL_HESS, tag
        assert relerr(nbs.segmpos_dparams_to_action_hess(v("segmpos"), v("dparams")), v("hess_cached")) < TOL_HESS, tag
        assert relerr(nbs.segmpos_params_to_action(v("segmpos"), v("params")), v("action")) < TOL_ACTION, tag
        assert relerr(nbs.segmpos_params_to_action_grad(v("segmpos"), v("params")), v("grad")) < TOL_GRAD, tag
        assert relerr(nbs.segmpos_to_params_T(v("segm")), v("params_T")) < TOL_TRANSFORM, tag
        assert relerr(nbs.segmpos_to_params(v("segmpos")), v("params_inv")) < TOL_TRANSFORM, tag


@pytest.mark.parametrize("name,nint", [("3B", 4096), ("3C", 3072), ("3C", 3 * 2 * 600)])
def test_hostsim_fused_radix16_sizes(name, nint, sim):
    """The fused CTA program at the bench sizes (radix-16/8 passes, swizzled buffer) against the oracle."""
    import os
    from conftest import ROOT
    from oracle.choreo_oracle import OracleSyst
    T = PlanTables.load(os.path.join(ROOT, "choreographies2_b200", "systems", f"{name}.npz"))
    nbs = NBodySystB200(T, nint, _lib_handle=sim)
    assert nbs.fused
    O = OracleSyst(T, nint, backend="fft")
    rng = np.random.default_rng(0)
    x, dx = rng.random(nbs.nparams), rng.random(nbs.nparams)
    act, grad = nbs.params_to_action_and_grad(x)
    seg = nbs.params_to_segmpos(x)
    assert relerr(act, O.params_to_action(x)) < TOL_ACTION
    assert relerr(grad, O.params_to_action_grad(x)) < TOL_GRAD
    assert relerr(seg, O.params_to_segmpos(x)) < TOL_TRANSFORM
    assert relerr(nbs.segmpos_dparams_to_action_hess(seg, dx), O.segmpos_dparams_to_action_hess(seg, dx)) < TOL_HESS


def test_host_layer_errors_and_batching(sim, golden_cache):
    g = golden_cache("3D")
    T = PlanTables.from_dict(g.tables_dict("base"))
    nint = max(g.variants["base"])
    nbs = NBodySystB200(T, nint, _lib_handle=sim)
    # illegal nint: same exception type and message as the reference setter (_NBodySyst.pyx:545-546)
    with pytest.raises(ValueError, match="should be divisible by"):
        nbs.nint = nint + 1
    assert nbs.nint == nint
    with pytest.raises(ValueError):
        nbs.params_to_action(np.zeros(nbs.nparams + 1))
    # batched call == loop of single calls, chunking does not change results
    rng = np.random.default_rng(0)
    x = rng.random((7, nbs.nparams))
    ref_act = np.array([nbs.params_to_action(x[i]) for i in range(7)])
    ref_grad = np.stack([nbs.params_to_action_grad(x[i]) for i in range(7)])
    for chunk in (0, 1, 3):
        nbs.set_chunk(chunk)
        act, grad = nbs.params_to_action_and_grad(x)
        assert np.array_equal(act, ref_act) and np.array_equal(grad, ref_grad)
    assert nbs.params_to_action_and_grad(x[:0])[1].shape == (0, nbs.nparams)  # empty batch
    assert nbs.kernel_launches > 0
    # nint_fac setter re-plans like the reference (find.py:314-317)
    nbs.nint_fac = 2 * nbs.nint_fac
    assert nbs.nint == 2 * nint and nbs.segm_size == 2 * nint // nbs.nint_min
    g3 = nbs.params_to_action_grad(rng.random(nbs.nparams))
    assert g3.shape == (nbs.nparams,) and np.isfinite(g3).all()


def test_unsupported_inputs_are_rejected(sim, golden_cache):
    g = golden_cache("3C")
    T = PlanTables.from_dict(g.tables_dict("base"))
    import dataclasses
    bad = dataclasses.replace(T, inter_law=7)
    with pytest.raises(ValueError, match="unsupported"):
        NBodySystB200(bad, 24, _lib_handle=sim)
    nbs = NBodySystB200(T, 24, _lib_handle=sim)
    with pytest.raises(ValueError, match="prime factor"):
        nbs.nint = 2 * 3 * 37  # FFT length 222 = 2*3*37
    class Fake:  # custom inter_law on the reference object
        inter_law_str = "def f(x): ..."
        inter_law_param_dict = None
    from choreographies2_b200.tables import _law_from_reference
    with pytest.raises(ValueError, match="Unsupported inter_law"):
        _law_from_reference(Fake())


def test_krylov_jacobian_glue(sim, golden_cache):
    """GetKrylovJacobian mirrors _NBodySyst.pyx:2004-2032: matvec = Hessian-vector with cached segmpos."""
    g = golden_cache("3C")
    nint = max(g.variants["base"])
    nbs = NBodySystB200(PlanTables.from_dict(g.tables_dict("base")), nint, _lib_handle=sim)
    v = lambda k: g.vec("base", nint, k)  # noqa: E731
    jac = nbs.GetKrylovJacobian(True, {"method": "gmres"})
    x = v("params")
    jac.setup(x.copy(), nbs.params_to_action_grad(x), nbs.params_to_action_grad)
    assert relerr(jac.matvec(v("dparams")), v("hess_cached")) < TOL_HESS
    assert relerr(jac.segmpos, v("segmpos")) < TOL_TRANSFORM


def test_hostsim_fused_v4_persistent_batch(sim):
    """Version 4 of the fused CTA program (chb_fused4.cuh) with more states than CTAs in the emulated grid: every CTA
    refills its buffers from the next state's rows; results must equal version 3 to rounding and the oracle to tolerance."""
    import os
    from conftest import ROOT
    from oracle.choreo_oracle import OracleSyst
    T = PlanTables.load(os.path.join(ROOT, "choreographies2_b200", "systems", "3B.npz"))
    nbs = NBodySystB200(T, 4096, _lib_handle=sim)
    assert nbs.fused_version == 4
    B = 150  # the emulated grid has 148 CTAs: states 148, 149 are second states
    rng = np.random.default_rng(5)
    x, dx = rng.random((B, nbs.nparams)), rng.random((B, nbs.nparams))
    a4, g4 = nbs.params_to_action_and_grad(x)
    seg = nbs.params_to_segmpos(x[-2:])
    h4 = nbs.segmpos_dparams_to_action_hess(seg, dx[-2:])
    nbs.set_fused(3)
    assert nbs.fused_version == 3
    a3, g3 = nbs.params_to_action_and_grad(x)
    h3 = nbs.segmpos_dparams_to_action_hess(seg, dx[-2:])
    assert relerr(a4, a3) < TOL_ACTION and relerr(g4, g3) < TOL_GRAD and relerr(h4, h3) < TOL_HESS
    O = OracleSyst(T, 4096, backend="fft")
    for b in (0, 148, 149):
        assert relerr(a4[b], O.params_to_action(x[b])) < TOL_ACTION
        assert relerr(g4[b], O.params_to_action_grad(x[b])) < TOL_GRAD
