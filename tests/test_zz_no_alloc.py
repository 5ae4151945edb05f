"""chb_params_to_action_hess_noalloc_dev: the Hessian-vector product for callers that cannot allocate device memory (no counterpart
in the reference).  The scratch is a device array built into the library.  Called
through the C ABI and compared with chb_params_to_action_hess_dev (caller's workspace) on two calls in a row: the second call finds
the first call's contents in the shared scratch.  This file sorts last on purpose."""
import ctypes as C
import os

import numpy as np
import pytest

from conftest import ROOT, TOL_HESS, relerr

pytestmark = pytest.mark.gpu


def _plan(name, nint, fused):
    from choreographies2_b200 import NBodySystB200, PlanTables
    nbs = NBodySystB200(PlanTables.load(os.path.join(ROOT, "choreographies2_b200", "systems", f"{name}.npz")), nint)
    nbs.set_fused(fused)
    return nbs


def _hess_with_workspace(nbs, x, dx):
    """The existing entry point through the C ABI: workspace sized by chb_workspace_bytes, allocated by the test."""
    import torch
    need = C.c_size_t(0)
    assert nbs._lib.chb_workspace_bytes(nbs._plan, x.shape[0], C.byref(need)) == 0
    ws = torch.empty(need.value, dtype=torch.uint8, device=x.device)
    out = torch.empty_like(x)
    st = C.c_void_p(torch.cuda.current_stream(x.device).cuda_stream)
    rc = nbs._lib.chb_params_to_action_hess_dev(nbs._plan, x.shape[0], C.c_void_p(x.data_ptr()), C.c_void_p(dx.data_ptr()),
                                                C.c_void_p(out.data_ptr()), C.c_void_p(ws.data_ptr()), need.value, st)
    assert rc == 0, nbs._lib.chb_last_error_string()
    return out


def _hess_noalloc(nbs, x, dx, out):
    import torch
    st = C.c_void_p(torch.cuda.current_stream(x.device).cuda_stream)
    return nbs._lib.chb_params_to_action_hess_noalloc_dev(nbs._plan, x.shape[0], C.c_void_p(x.data_ptr()), C.c_void_p(dx.data_ptr()),
                                                          C.c_void_p(out.data_ptr()), st)


@pytest.mark.parametrize("name,nint,fused,batch", [("3C", 768, True, 37), ("3C", 768, False, 37), ("20B", 240, True, 9), ("5Dq", 640, True, 12)])
def test_noalloc_hessian_vector_matches_workspace_variant_twice(name, nint, fused, batch):
    import torch
    nbs = _plan(name, nint, fused)
    size = C.c_size_t(0)
    assert nbs._lib.chb_noalloc_scratch_bytes(C.byref(size)) == 0 and size.value == nbs.NOALLOC_SCRATCH_BYTES == 32 << 20
    rng = np.random.default_rng(3)
    out = torch.full((batch, nbs.nparams), float("nan"), dtype=torch.float64, device="cuda")
    for call in range(2):  # the second call runs on the scratch the first one left behind, with other inputs
        x = torch.from_numpy(rng.random((batch, nbs.nparams))).cuda()
        dx = torch.from_numpy(rng.random((batch, nbs.nparams))).cuda()
        want = _hess_with_workspace(nbs, x, dx)
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated()
        assert _hess_noalloc(nbs, x, dx, out) == 0, nbs._lib.chb_last_error_string()
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated() == before  # nothing allocated through torch either
        assert torch.isfinite(out).all()
        assert relerr(out.cpu().numpy(), want.cpu().numpy()) < TOL_HESS, (name, nint, fused, call)
    # the host-layer method is the same call
    out2 = torch.empty_like(out)
    nbs.params_to_action_hess_noalloc(x, dx, out2)
    assert torch.equal(out2, out)


def test_noalloc_rejects_batches_beyond_the_built_in_scratch():
    import torch
    from choreographies2_b200._lib import ChoreoB200Error
    nbs = _plan("3C", 3072, False)
    batch = 2048
    nbs.set_chunk(batch)  # one chunk: the scratch need follows the whole batch
    need = C.c_size_t(0)
    assert nbs._lib.chb_workspace_bytes(nbs._plan, batch, C.byref(need)) == 0 and need.value > 32 << 20
    x = torch.from_numpy(np.random.default_rng(5).random((batch, nbs.nparams))).cuda()
    out = torch.full_like(x, 7.0)
    rc = _hess_noalloc(nbs, x, x, out)
    assert rc == 6  # CHB_ERR_WORKSPACE
    msg = nbs._lib.chb_last_error_string().decode()
    assert "CHB_NOALLOC_SCRATCH_BYTES" in msg and str(32 << 20) in msg
    torch.cuda.synchronize()
    assert bool((out == 7.0).all())  # nothing was launched
    with pytest.raises(ChoreoB200Error, match="CHB_ERR_WORKSPACE"):
        nbs.params_to_action_hess_noalloc(x, x, out)
    nbs.set_chunk(64)  # smaller chunks of states: the same batch fits
    need2 = C.c_size_t(0)
    assert nbs._lib.chb_workspace_bytes(nbs._plan, batch, C.byref(need2)) == 0 and need2.value <= 32 << 20
    assert _hess_noalloc(nbs, x, x, out) == 0, nbs._lib.chb_last_error_string()
    torch.cuda.synchronize()
    assert torch.isfinite(out).all() and not bool((out == 7.0).all())
