"""Drop-in use of the B200 path under the reference's own search driver (``choreo/find.py``, ``choreo/run.py``).

``NBodySyst`` is a ``@cython.final cdef class`` (_NBodySyst.pyx:153-154): it cannot be subclassed, and ``Find_Choreo`` builds its
own instance (find.py:94).  ``install(choreo)`` therefore replaces the name ``choreo.NBodySyst`` by ``NBodySystDropIn``: a thin
object that builds the reference ``NBodySyst`` as before (symmetry setup, descriptors, plotting stay the reference's), reads its
tables once, and routes the numeric methods of the action path (SURVEY.md 8(a) row a15 and the post-solve reductions of 8(f)
rank 2) to ``NBodySystB200`` -- one state per call, NumPy in / NumPy out, exactly as ``find.py`` and ``nonlin_solve_pp`` call
them.  Every other attribute falls through to the reference object.

    import choreo, choreographies2_b200.dropin as dropin
    with dropin.installed(choreo):
        choreo.find.ChoreoReadDictAndFind(workspace_folder)

A batch of one state per call leaves the GPU idle most of the time (launch latency); the batched driver
(``refine.find_batched``) is the productive use.  This module exists to show -- and test -- that the boundary is the
reference's own.
"""
from __future__ import annotations

import contextlib
from typing import Any

from .nbodysyst import NBodySystB200
from .tables import extract_tables

# methods of NBodySyst served by the library (reference names, _NBodySyst.pyx:2004-3110, 3607-3727, 2125-2192, 2542-2560, 3968-4009)
ROUTED = (
    "params_to_action", "params_to_action_grad", "params_to_action_hess", "segmpos_params_to_action", "segmpos_params_to_action_grad",
    "segmpos_dparams_to_action_hess", "params_to_segmpos", "params_to_segmvel", "segmpos_to_params", "segmpos_to_params_T",
    "params_to_kin_nrg", "params_to_kin_nrg_grad", "params_to_pot_nrg", "params_to_pot_nrg_grad", "params_to_pot_nrg_hess",
    "params_changevar", "params_resize", "segmpos_to_hash", "DetectEscape", "ComputeCenterOfMass", "GetKrylovJacobian",
)


class NBodySystDropIn:
    """Reference ``NBodySyst`` with the action path on the B200.  ``_lib_handle`` / ``_device`` are class-level knobs the tests
    use to run the same object over the CPU emulation of the kernels."""
    _reference_class: Any = None
    _lib_handle: Any = None
    _device: int = 0
    calls: dict = {}

    def __init__(self, *args, **kwargs):
        object.__setattr__(self, "_ref", type(self)._reference_class(*args, **kwargs))
        object.__setattr__(self, "_gpu", None)

    def _b200(self) -> NBodySystB200:
        ref = object.__getattribute__(self, "_ref")
        gpu = object.__getattribute__(self, "_gpu")
        if gpu is None:
            gpu = NBodySystB200(extract_tables(ref), int(ref.nint), device=type(self)._device, _lib_handle=type(self)._lib_handle)
            object.__setattr__(self, "_gpu", gpu)
        elif gpu.nint != int(ref.nint):
            gpu.nint = int(ref.nint)  # the reference's nint / nint_fac setters re-plan (find.py:101, 314-317, 429)
        return gpu

    def __getattr__(self, name):
        if name in ROUTED:
            fn = getattr(self._b200(), name)
            calls = type(self).calls

            def routed(*a, **k):
                calls[name] = calls.get(name, 0) + 1
                out = fn(*a, **k)
                if name == "GetKrylovJacobian":
                    out.NBS = self  # the Jacobian's matvec / update go through this object too (and are counted)
                return out
            return routed
        return getattr(object.__getattribute__(self, "_ref"), name)

    def __setattr__(self, name, value):
        # properties of the reference object (nint, nint_fac, fft_backend, fftw_* ...) keep their setters
        setattr(object.__getattribute__(self, "_ref"), name, value)


def install(choreo_module, lib_handle=None, device: int = 0):
    """Replace ``choreo.NBodySyst`` (and the name inside ``choreo.cython``) by the drop-in; returns the original class."""
    original = choreo_module.NBodySyst
    if original is NBodySystDropIn:
        return NBodySystDropIn._reference_class
    NBodySystDropIn._reference_class = original
    NBodySystDropIn._lib_handle = lib_handle
    NBodySystDropIn._device = device
    NBodySystDropIn.calls = {}
    choreo_module.NBodySyst = NBodySystDropIn
    return original


def uninstall(choreo_module, original) -> None:
    choreo_module.NBodySyst = original


@contextlib.contextmanager
def installed(choreo_module, lib_handle=None, device: int = 0):
    original = install(choreo_module, lib_handle, device)
    try:
        yield NBodySystDropIn
    finally:
        uninstall(choreo_module, original)
