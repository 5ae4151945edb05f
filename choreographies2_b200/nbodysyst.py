"""Host-side mirror of the reference ``NBodySyst`` API for the action hot path, backed by the B200 library.

Same method names, argument meaning and error behaviour as the reference's Python-visible methods
(choreo/cython/_NBodySyst.pyx:2519-3110, 3607-3727), so that
``find.py`` / ``scipy_plus/nonlin.py`` can drive it unchanged for one state, plus ``*_batched`` variants that
evaluate many independent states per call (leading batch dimension).

Array carriers:
  * NumPy arrays (host): staged through the library (``chb_run_host``: H2D, kernels, D2H, synchronous);
  * torch CUDA tensors (optional): passed by ``data_ptr()`` to the ``*_dev`` entry points on torch's current
    stream; results are torch tensors on the same device.  torch is only an array carrier here.

There is no CPU fallback: constructing the object without the CUDA library / an sm_100 device raises.
"""
from __future__ import annotations

import ctypes as C
import types
from typing import Any, Optional

import numpy as np

from . import _lib as L
from .tables import PlanTables, extract_tables


def _is_torch(x: Any) -> bool:
    return type(x).__module__.startswith("torch")


class NBodySystB200:
    def __init__(self, tables: PlanTables, nint: Optional[int] = None, device: int = 0, _lib_handle=None):
        self._lib = _lib_handle if _lib_handle is not None else L.load()
        self.tables = tables
        self.device = int(device)
        self._plan = C.c_void_p()
        struct, keep = L.tables_struct(tables)
        L.check(self._lib, self._lib.chb_plan_create(C.byref(struct), self.device, C.byref(self._plan)))
        del keep
        self._ws = None  # torch workspace tensor (device path)
        self._nint = 0
        self.nint = int(nint) if nint is not None else 2 * tables.nint_min  # reference: nint_fac = 1 (:895)

    # ------------------------------------------------------------------ construction helpers
    @classmethod
    def from_reference(cls, NBS, device: int = 0, _lib_handle=None) -> "NBodySystB200":
        """Build from a live reference ``choreo.NBodySyst`` (tables are read from its public attributes)."""
        return cls(extract_tables(NBS), nint=int(NBS.nint), device=device, _lib_handle=_lib_handle)

    @classmethod
    def from_file(cls, path: str, nint: Optional[int] = None, device: int = 0) -> "NBodySystB200":
        return cls(PlanTables.load(path), nint=nint, device=device)

    def __del__(self):
        try:
            if getattr(self, "_plan", None) is not None and self._plan.value:
                self._lib.chb_plan_destroy(self._plan)
                self._plan = C.c_void_p()
        except Exception:
            pass

    # ------------------------------------------------------------------ properties (reference names)
    def _q(self, what: int) -> int:
        v = C.c_int64()
        L.check(self._lib, self._lib.chb_plan_query(self._plan, what, C.byref(v)))
        return int(v.value)

    geodim = property(lambda self: self.tables.geodim)
    nbody = property(lambda self: self.tables.nbody)
    nloop = property(lambda self: self.tables.nloop)
    nsegm = property(lambda self: self.tables.nsegm)
    nint_min = property(lambda self: self.tables.nint_min)
    nbin_segm_unique = property(lambda self: self.tables.nbin)
    GreaterNStore = property(lambda self: self.tables.greater_nstore)

    @property
    def nint(self) -> int:
        return self._nint

    @nint.setter
    def nint(self, nint_in: int) -> None:
        # reference setter: ValueError when nint % (2*nint_min) != 0 (:545-546); re-plans all buffers (:589-590)
        L.check(self._lib, self._lib.chb_plan_set_nint(self._plan, int(nint_in)))
        self._nint = int(nint_in)
        self.nparams = self._q(L.Q_NPARAMS)
        self.segm_size = self._q(L.Q_SEGM_SIZE)
        self.segm_store = self._q(L.Q_SEGM_STORE)
        self.ncoeffs = self._nint // 2 + 1
        self._segsz = self._q(L.Q_SEGMPOS_SIZE)
        self.nparams_incl_o = self._q(L.Q_NPARAMS_INCL_O)
        self._ws = None

    @property
    def nint_fac(self) -> int:
        return self._nint // (2 * self.nint_min)

    @nint_fac.setter
    def nint_fac(self, fac: int) -> None:
        self.nint = 2 * self.nint_min * int(fac)

    @property
    def segmpos_shape(self):
        return (self.nsegm, self.segm_store, self.geodim)

    @property
    def kernel_launches(self) -> int:
        return self._q(L.Q_LAUNCHES)

    @property
    def fused(self) -> bool:
        return bool(self._q(L.Q_FUSED))

    @property
    def fused_version(self) -> int:
        """0 = staged kernels, 3 / 4 = version of the fused kernel serving this nint level."""
        return self._q(L.Q_FUSED)

    @property
    def staged_kernels(self):
        """(transform kind, pair kernel) of the staged path: (0 generic | sum of the available residue splits R of
        ``chb_big.cuh``, e.g. 14 = 2 + 4 + 8; 0 colour rounds | 1 dense | 2 tiled)."""
        return self._q(L.Q_STAGED_TRANSFORM), self._q(L.Q_PAIR_KERNEL)

    def set_chunk(self, chunk: int) -> None:
        L.check(self._lib, self._lib.chb_plan_set_chunk(self._plan, int(chunk)))
        self._ws = None

    def set_fused(self, enable) -> None:
        """False/0: generic staged kernels only; True/2: fused kernel (default); 1: fused kernel with separate
        load/store phases (kept for A/B measurements)."""
        level = 4 if enable is True else int(enable)
        L.check(self._lib, self._lib.chb_plan_set_fused(self._plan, level))

    def set_transform_split(self, split: int) -> None:
        """0: residue split of the staged transforms chosen per launch (default); 2, 4, 8: that split whenever the level has it."""
        L.check(self._lib, self._lib.chb_plan_set_split(self._plan, int(split)))

    # ------------------------------------------------------------------ plumbing
    def _check(self, arr, last_shape, what: str):
        shp = tuple(arr.shape)
        n = len(last_shape)
        if shp[-n:] != tuple(last_shape):
            raise ValueError(f"{what}: expected trailing shape {tuple(last_shape)}, got {shp}")
        batch = int(np.prod(shp[:-n], dtype=np.int64)) if len(shp) > n else 1
        return batch, shp[:-n]

    def _out(self, given, shape, what):
        if given is None:
            return np.empty(shape, dtype=np.float64)
        if (not isinstance(given, np.ndarray) or given.dtype != np.float64 or not given.flags.c_contiguous
                or tuple(given.shape) != tuple(shape)):
            raise ValueError(f"{what}: expected a C-contiguous float64 array of shape {tuple(shape)}")
        return given

    def pinned_empty(self, shape) -> np.ndarray:
        """Page-locked float64 NumPy array (``chb_malloc_pinned``): host buffers of this kind let the host-pointer
        operators overlap H2D, kernels and D2H.  Freed when the array (and every view of it) is collected."""
        import weakref
        shape = tuple(int(v) for v in (shape if hasattr(shape, "__len__") else (shape,)))
        n = int(np.prod(shape, dtype=np.int64))
        ptr = C.c_void_p()
        L.check(self._lib, self._lib.chb_malloc_pinned(max(n, 1) * 8, C.byref(ptr)))
        buf = (C.c_double * max(n, 1)).from_address(ptr.value)
        arr = np.frombuffer(buf, dtype=np.float64, count=n).reshape(shape)
        weakref.finalize(buf, self._lib.chb_free_pinned, C.c_void_p(ptr.value))
        return arr

    def _host(self, op, in0, in0_shape, in1=None, in1_shape=None, out0_shape=None, out1_shape=None, scalar=0.0,
              want_out0=True, out0=None, out1=None):
        a0 = np.ascontiguousarray(in0, dtype=np.float64)
        batch, lead = self._check(a0, in0_shape, "input")
        a1 = None
        if in1 is not None:
            a1 = np.ascontiguousarray(in1, dtype=np.float64)
            b1, _ = self._check(a1, in1_shape, "second input")
            if b1 != batch:
                raise ValueError("batch sizes of the two inputs differ")
        o0 = self._out(out0, lead + tuple(out0_shape), "out0") if (out0_shape is not None and want_out0) else None
        o1 = self._out(out1, lead + tuple(out1_shape), "out1") if out1_shape is not None else None
        vp = lambda a: a.ctypes.data_as(C.c_void_p) if a is not None else None  # noqa: E731
        L.check(self._lib, self._lib.chb_run_host(self._plan, op, batch, vp(a0), vp(a1), float(scalar), vp(o0), vp(o1)))
        return o0, o1

    def _workspace(self, batch: int, like):
        import torch
        need = C.c_size_t()
        L.check(self._lib, self._lib.chb_workspace_bytes(self._plan, batch, C.byref(need)))
        if self._ws is None or self._ws.numel() < need.value or self._ws.device != like.device:
            self._ws = torch.empty(int(need.value), dtype=torch.uint8, device=like.device)
        return self._ws

    def _dev(self, fname, batch, ptr_args, like):
        """Call a *_dev entry point: ptr_args is the list between (plan, batch) and (ws, ws_bytes, stream)."""
        import torch
        ws = self._workspace(batch, like)
        stream = torch.cuda.current_stream(like.device).cuda_stream
        fn = getattr(self._lib, fname)
        conv = [a if isinstance(a, float) else (C.c_void_p(a.data_ptr()) if a is not None else None) for a in ptr_args]
        L.check(self._lib, fn(self._plan, batch, *conv, C.c_void_p(ws.data_ptr()), ws.numel(), C.c_void_p(stream)))

    def _tprep(self, t, last_shape, what):
        import torch
        if t.dtype != torch.float64 or not t.is_cuda:
            raise ValueError(f"{what}: torch tensors must be float64 CUDA tensors")
        if t.device.index != self.device:
            raise ValueError(f"{what}: tensor on {t.device}, plan on cuda:{self.device}")
        t = t.contiguous()
        batch, lead = self._check(t, last_shape, what)
        return t, batch, lead

    def _arr_call(self, fname, ins, outs, make_args):
        """Call a ``*_dev`` entry point that takes no workspace.  ``ins``: input arrays (all NumPy or all torch CUDA tensors,
        contiguous); ``outs``: list of ``(shape, dtype)`` of the results; ``make_args(in_ptrs, out_ptrs)`` builds the arguments
        between the plan and the stream.  torch tensors: called on the current stream; NumPy arrays: addressed directly when the
        library's ``*_dev`` entry points take host pointers, else staged through device buffers of the library
        (``chb_malloc_device`` / ``chb_memcpy_*``; these operators are post-solve reductions, not the hot path)."""
        fn = getattr(self._lib, fname)
        if _is_torch(ins[0]):
            import torch
            tdt = {np.float64: torch.float64, np.int32: torch.int32}
            res = [torch.empty(shp, dtype=tdt[dt], device=ins[0].device) for shp, dt in outs]
            stream = C.c_void_p(torch.cuda.current_stream(ins[0].device).cuda_stream)
            vp = lambda t: C.c_void_p(t.data_ptr())  # noqa: E731
            L.check(self._lib, fn(self._plan, *make_args([vp(a) for a in ins], [vp(r) for r in res]), stream))
            return res
        res = [np.empty(shp, dtype=dt) for shp, dt in outs]
        if self._q(L.Q_HOST_POINTERS):
            vp = lambda a: C.c_void_p(a.ctypes.data)  # noqa: E731
            L.check(self._lib, fn(self._plan, *make_args([vp(a) for a in ins], [vp(r) for r in res]), None))
            return res
        bufs = []
        try:
            def dev(nbytes):
                ptr = C.c_void_p()
                L.check(self._lib, self._lib.chb_malloc_device(self._plan, max(int(nbytes), 16), C.byref(ptr)))
                bufs.append(ptr)
                return ptr
            din = [dev(a.nbytes) for a in ins]
            dout = [dev(r.nbytes) for r in res]
            for a, d in zip(ins, din):
                L.check(self._lib, self._lib.chb_memcpy_h2d(self._plan, d, C.c_void_p(a.ctypes.data), a.nbytes, None))
            L.check(self._lib, fn(self._plan, *make_args(din, dout), None))
            for r, d in zip(res, dout):
                L.check(self._lib, self._lib.chb_memcpy_d2h(self._plan, C.c_void_p(r.ctypes.data), d, r.nbytes, None))
            L.check(self._lib, self._lib.chb_stream_synchronize(self._plan, None))
        finally:
            for ptr in bufs:
                self._lib.chb_free_device(self._plan, ptr)
        return res

    def _prep(self, a, last_shape, what):
        """(contiguous float64 array, batch, leading shape) for a NumPy array or a torch CUDA tensor."""
        if _is_torch(a):
            return self._tprep(a, last_shape, what)
        a = np.ascontiguousarray(a, dtype=np.float64)
        batch, lead = self._check(a, last_shape, what)
        return a, batch, lead

    def params_to_segmvel(self, params_mom_buf):
        """_NBodySyst.pyx:3638-3666 (cdef :6050-6118): segment velocities, shape of ``segmpos``."""
        if _is_torch(params_mom_buf):
            import torch
            x, b, lead = self._tprep(params_mom_buf, (self.nparams,), "params")
            out = torch.empty(lead + self.segmpos_shape, dtype=torch.float64, device=x.device)
            if b > 0:
                # a call with an empty batch creates the velocity plan, so that the workspace query covers it
                L.check(self._lib, self._lib.chb_params_to_segmvel_dev(self._plan, 0, None, None, None, 0, None))
                self._ws = None
                self._dev("chb_params_to_segmvel_dev", b, [x, out], x)
            return out
        return self._host(L.OP_PARAMS_TO_SEGMVEL, params_mom_buf, (self.nparams,), out0_shape=self.segmpos_shape)[0]

    def params_changevar(self, params_buf_in, inv: bool = False, transpose: bool = False):
        """_NBodySyst.pyx:2330-2410: the diagonal change of variables between the "momentum" parameters (``nparams``) and the
        position Fourier parameters (``nparams_incl_o``, zero-padded), its transpose, inverse and inverse transpose."""
        to_pos = (not inv and not transpose) or (inv and transpose)
        n_in, n_out = (self.nparams, self.nparams_incl_o) if to_pos else (self.nparams_incl_o, self.nparams)
        x, b, lead = self._prep(params_buf_in, (n_in,), "params_buf_in")
        out, = self._arr_call("chb_params_changevar_dev", [x], [(lead + (n_out,), np.float64)],
                              lambda i, o: (b, i[0], int(bool(inv)), int(bool(transpose)), o[0]))
        return out

    # ------------------------------------------------------------------ transforms
    def params_to_segmpos(self, params_mom_buf):
        """_NBodySyst.pyx:3607-3635."""
        if _is_torch(params_mom_buf):
            import torch
            x, b, lead = self._tprep(params_mom_buf, (self.nparams,), "params")
            out = torch.empty(lead + self.segmpos_shape, dtype=torch.float64, device=x.device)
            self._dev("chb_params_to_segmpos_dev", b, [x, out], x)
            return out
        return self._host(L.OP_PARAMS_TO_SEGMPOS, params_mom_buf, (self.nparams,), out0_shape=self.segmpos_shape)[0]

    def segmpos_to_params(self, segmpos):
        """_NBodySyst.pyx:3669-3696."""
        if _is_torch(segmpos):
            import torch
            x, b, lead = self._tprep(segmpos, self.segmpos_shape, "segmpos")
            out = torch.empty(lead + (self.nparams,), dtype=torch.float64, device=x.device)
            self._dev("chb_segmpos_to_params_dev", b, [x, out], x)
            return out
        return self._host(L.OP_SEGMPOS_TO_PARAMS, segmpos, self.segmpos_shape, out0_shape=(self.nparams,))[0]

    def segmpos_to_params_T(self, segmpos):
        """_NBodySyst.pyx:3699-3727."""
        if _is_torch(segmpos):
            import torch
            x, b, lead = self._tprep(segmpos, self.segmpos_shape, "segmpos")
            out = torch.empty(lead + (self.nparams,), dtype=torch.float64, device=x.device)
            self._dev("chb_segmpos_to_params_T_dev", b, [x, out], x)
            return out
        return self._host(L.OP_SEGMPOS_TO_PARAMS_T, segmpos, self.segmpos_shape, out0_shape=(self.nparams,))[0]

    # ------------------------------------------------------------------ energies
    @staticmethod
    def _scalar(arr, lead):
        return float(arr.reshape(-1)[0]) if lead == () else arr.reshape(lead)

    def params_to_kin_nrg(self, params_mom_buf):
        """_NBodySyst.pyx:2519-2525."""
        if _is_torch(params_mom_buf):
            import torch
            x, b, lead = self._tprep(params_mom_buf, (self.nparams,), "params")
            out = torch.empty(lead, dtype=torch.float64, device=x.device)
            self._dev("chb_params_to_kin_nrg_dev", b, [x, out], x)
            return out
        a = np.asarray(params_mom_buf)
        o, _ = self._host(L.OP_KIN, a, (self.nparams,), out0_shape=())
        return self._scalar(o, a.shape[:-1])

    def params_to_kin_nrg_grad(self, params_mom_buf):
        """_NBodySyst.pyx:2527-2540: the parameters themselves on every k>0 entry, 0 elsewhere."""
        a = np.ascontiguousarray(params_mom_buf, dtype=np.float64)
        lvl = self.tables.level(self._nint)
        mask = np.ones(self.nparams, dtype=bool)
        for il in range(self.nloop):
            ncor, _ = self.tables.ncor_nco_in(il)
            mask[lvl.mom_off[il]:lvl.mom_off[il] + ncor] = False
        return a * mask

    def segmpos_to_pot_nrg(self, segmpos):
        """cdef segm_pos_to_pot_nrg, _NBodySyst.pyx:6372-6469."""
        if _is_torch(segmpos):
            import torch
            x, b, lead = self._tprep(segmpos, self.segmpos_shape, "segmpos")
            out = torch.empty(lead, dtype=torch.float64, device=x.device)
            self._dev("chb_segmpos_to_pot_nrg_dev", b, [x, out], x)
            return out
        a = np.asarray(segmpos)
        o, _ = self._host(L.OP_POT, a, self.segmpos_shape, out0_shape=())
        return self._scalar(o, a.shape[:-3])

    # default exponents of the reference (``default_Hash_exp``, _NBodySyst.pyx:135)
    Hash_exp = np.array([-0.1, -0.2, -0.3, -0.4, -0.5, -0.6, -0.7, -0.8, -0.9])

    def segmpos_to_hash(self, segmpos):
        """_NBodySyst.pyx:2542-2560 (cdef segm_pos_to_hash :6271-6369): ``len(Hash_exp)`` weighted power sums of the pair
        distances, the orbit fingerprint used to filter duplicates (find.py:285-308, 905-935); all exponents in ONE pass over
        the pair distances (``chb_segmpos_to_hash_multi_dev``).  Returns ``[..., nexp]``."""
        x, b, lead = self._prep(segmpos, self.segmpos_shape, "segmpos")
        exps = np.ascontiguousarray(self.Hash_exp, dtype=np.float64)
        nexp = int(exps.shape[0])
        if b == 0:
            return (x.new_empty if _is_torch(x) else np.empty)(lead + (nexp,))
        out, = self._arr_call("chb_segmpos_to_hash_multi_dev", [x], [(lead + (nexp,), np.float64)],
                              lambda i, o: (b, i[0], nexp, C.c_void_p(exps.ctypes.data), o[0]))
        return out

    # ------------------------------------------------------------------ ODE right-hand sides
    def _forces(self, pos_flat, dpos_flat=None):
        vs = self.nsegm * self.geodim
        if _is_torch(pos_flat):
            import torch
            x, nvec, _ = self._tprep(pos_flat, (vs,), "pos_flat")
            stream = C.c_void_p(torch.cuda.current_stream(x.device).cuda_stream)
            vp = lambda a: C.c_void_p(a.data_ptr())  # noqa: E731
            if dpos_flat is None:
                out = torch.empty_like(x)
                L.check(self._lib, self._lib.chb_compute_forces_dev(self._plan, nvec, vp(x), vp(out), stream))
                return out
            if not _is_torch(dpos_flat) or dpos_flat.dim() < 2 or dpos_flat.shape[-2] != vs:
                raise ValueError(f"dpos_flat: expected a tensor of shape [..., {vs}, grad_ndof]")
            ndof = int(dpos_flat.shape[-1])
            d, nd, _ = self._tprep(dpos_flat, (vs, ndof), "dpos_flat")
            if nd != nvec:
                raise ValueError("pos_flat and dpos_flat hold different numbers of vectors")
            out = torch.empty_like(d)
            L.check(self._lib, self._lib.chb_compute_grad_forces_dev(self._plan, nvec, ndof, vp(x), vp(d), vp(out), stream))
            return out
        x = np.ascontiguousarray(pos_flat, dtype=np.float64)
        nvec, _ = self._check(x, (vs,), "pos_flat")
        vp = lambda a: a.ctypes.data_as(C.c_void_p) if a is not None else None  # noqa: E731
        if dpos_flat is None:
            out = np.empty_like(x)
            L.check(self._lib, self._lib.chb_compute_forces_host(self._plan, nvec, 0, vp(x), None, vp(out), None))
            return out
        d = np.ascontiguousarray(dpos_flat, dtype=np.float64)
        if d.ndim < 2 or d.shape[-2] != vs:
            raise ValueError(f"dpos_flat: expected an array of shape [..., {vs}, grad_ndof]")
        ndof = int(d.shape[-1])
        nd, _ = self._check(d, (vs, ndof), "dpos_flat")
        if nd != nvec:
            raise ValueError("pos_flat and dpos_flat hold different numbers of vectors")
        out = np.empty_like(d)
        L.check(self._lib, self._lib.chb_compute_forces_host(self._plan, nvec, ndof, vp(x), vp(d), None, vp(out)))
        return out

    def Compute_forces(self, t, pos_flat):
        """_NBodySyst.pyx:4154-4171 (cdef 7531-7638): forces on the segments at one phase-space point, ``[nsegm*geodim]``
        (``t`` is unused, as in the reference: the system is autonomous)."""
        return self._forces(pos_flat)

    def Compute_forces_vectorized(self, t, pos_flat):
        """_NBodySyst.pyx:4228-4246: ``pos_flat`` is ``[nvec, nsegm*geodim]`` (any leading shape here); one thread per
        vector on the device (``chb_compute_forces_dev``)."""
        return self._forces(pos_flat)

    def Compute_grad_forces(self, t, pos_flat, dpos_flat):
        """_NBodySyst.pyx:4173-4191 (cdef 7641-7790): tangent forces, ``dpos_flat`` and the result ``[nsegm*geodim, grad_ndof]``."""
        return self._forces(pos_flat, dpos_flat)

    def Compute_grad_forces_vectorized(self, t, pos_flat, dpos_flat):
        """_NBodySyst.pyx:4249-4268: ``[nvec, nsegm*geodim]`` positions, ``[nvec, nsegm*geodim, grad_ndof]`` tangents."""
        return self._forces(pos_flat, dpos_flat)

    def _velocities(self, mom_flat, grad_mom_flat=None):
        vs = self.nsegm * self.geodim
        if grad_mom_flat is None:
            x, nvec, lead = self._prep(mom_flat, (vs,), "mom_flat")
            if nvec == 0:
                return x
            out, = self._arr_call("chb_compute_velocities_dev", [x], [(tuple(x.shape), np.float64)], lambda i, o: (nvec, 0, i[0], o[0]))
            return out
        if len(grad_mom_flat.shape) < 2 or grad_mom_flat.shape[-2] != vs:
            raise ValueError(f"grad_mom_flat: expected shape [..., {vs}, grad_ndof]")
        ndof = int(grad_mom_flat.shape[-1])
        g, nvec, lead = self._prep(grad_mom_flat, (vs, ndof), "grad_mom_flat")
        if nvec == 0 or ndof == 0:
            return g
        out, = self._arr_call("chb_compute_velocities_dev", [g], [(tuple(g.shape), np.float64)], lambda i, o: (nvec, ndof, i[0], o[0]))
        return out

    def Compute_velocities(self, t, mom_flat):
        """_NBodySyst.pyx:4081-4094 (cdef 8106-8133): velocities of the segments from their momenta, ``mom / mass``."""
        return self._velocities(mom_flat)

    def Compute_velocities_vectorized(self, t, mom_flat):
        """_NBodySyst.pyx:4116-4131: ``[nvec, nsegm*geodim]``."""
        return self._velocities(mom_flat)

    def Compute_grad_velocities(self, t, mom_flat, grad_mom_flat):
        """_NBodySyst.pyx:4098-4113 (cdef 8135-8160): tangent of the above, ``[nsegm*geodim, grad_ndof]``."""
        return self._velocities(mom_flat, grad_mom_flat)

    def Compute_grad_velocities_vectorized(self, t, mom_flat, grad_mom_flat):
        """_NBodySyst.pyx:4133-4150: ``[nvec, nsegm*geodim, grad_ndof]``."""
        return self._velocities(mom_flat, grad_mom_flat)

    def Compute_init_pos_mom(self, params_mom_buf):
        """_NBodySyst.pyx:4060-4077: initial positions and momenta (first stored sample of every segment) of the reduced ODE."""
        seg = self.params_to_segmpos(params_mom_buf)
        vel = self.params_to_segmvel(params_mom_buf)
        mass = self.tables.loopmass[self.tables.bodyloop[self.tables.gensegm_to_body]]
        if _is_torch(seg):
            import torch
            mass = torch.as_tensor(mass, device=seg.device)
        xo = seg[..., :, 0, :].reshape(tuple(seg.shape[:-3]) + (-1,))
        po = (vel[..., :, 0, :] * mass[:, None]).reshape(tuple(seg.shape[:-3]) + (-1,))
        return xo, po

    def Get_ODE_def(self, params_mom_buf, vector_calls: bool = True, LowLevel: bool = False, NoSymIfPossible: bool = True,
                    grad: bool = False):
        """_NBodySyst.pyx:4310-4444: the reduced ODE of one segment period as a dict for the reference's symplectic integrators
        (``scipy_plus.ODE.SymplecticIVP``): Python callables backed by the library's kernels.  ``LowLevel`` C callbacks cannot
        reach the device and are rejected."""
        if LowLevel:
            raise ValueError("Get_ODE_def: LowLevelCallable C callbacks are CPU function pointers; use LowLevel=False")
        xo, po = self.Compute_init_pos_mom(params_mom_buf)
        d = {"t_span": (0., 1. / self.nint_min), "x0": xo, "v0": po, "vector_calls": vector_calls}
        if vector_calls:
            d["fun"], d["gun"] = self.Compute_velocities_vectorized, self.Compute_forces_vectorized
            if grad:
                d["grad_fun"], d["grad_gun"] = self.Compute_grad_velocities_vectorized, self.Compute_grad_forces_vectorized
        else:
            d["fun"], d["gun"] = self.Compute_velocities, self.Compute_forces
            if grad:
                d["grad_fun"], d["grad_gun"] = self.Compute_grad_velocities, self.Compute_grad_forces
        return d

    def params_resize(self, params_buf_in, nint_fac: int = 1):
        """_NBodySyst.pyx:2413-2465: the parameters of this ``nint`` level re-expressed at ``nint = 2 nint_min nint_fac``
        (batched over leading dimensions)."""
        return self.tables.params_resize(params_buf_in, self.nint, 2 * self.tables.nint_min * int(nint_fac))

    # ------------------------------------------------------------------ Krylov helper of the batched Newton driver
    def arnoldi_step(self, V, j: int, w, skip, h) -> bool:
        """One Arnoldi step of the batched GMRES for all states in one launch (``chb_arnoldi_step_dev``): orthogonalises
        ``w - V[j]`` against ``V[0..j]``, writes ``V[j+1]`` and ``h[:, :j+2]`` (see include/choreo_b200.h).  ``V`` is
        ``[>= j+2, batch, n]``, ``w`` ``[batch, n]``, ``skip`` ``[batch]`` bool or None, ``h`` ``[batch, ldh]``.  Returns False
        when the arrays are not ones the library can address (NumPy arrays with the CUDA library): the caller then runs its
        own array-level Gram-Schmidt."""
        torch_in = _is_torch(V)
        if not self._direct_ok(V):
            return False
        batch, n = int(w.shape[0]), int(w.shape[1])
        if j >= 62:
            return False
        if tuple(V.shape[1:]) != (batch, n) or V.shape[0] < j + 2 or tuple(h.shape[:1]) != (batch,) or h.shape[1] < j + 2:
            raise ValueError("arnoldi_step: inconsistent shapes")
        if torch_in:
            import torch
            for t in (V, w, h):
                if t.dtype != torch.float64 or not t.is_cuda or not t.is_contiguous():
                    raise ValueError("arnoldi_step: contiguous float64 CUDA tensors expected")
            sk = None
            if skip is not None:
                sk = skip.to(torch.uint8) if skip.dtype != torch.uint8 else skip
                sk = sk.contiguous()
            stream = torch.cuda.current_stream(V.device).cuda_stream
            L.check(self._lib, self._lib.chb_arnoldi_step_dev(
                self._plan, batch, n, int(j), int(h.shape[1]), C.c_void_p(V.data_ptr()), C.c_void_p(w.data_ptr()),
                C.c_void_p(sk.data_ptr()) if sk is not None else None, C.c_void_p(h.data_ptr()), C.c_void_p(stream)))
            return True
        for t in (V, w, h):
            if t.dtype != np.float64 or not t.flags.c_contiguous:
                raise ValueError("arnoldi_step: C-contiguous float64 arrays expected")
        sk = np.ascontiguousarray(skip, dtype=np.uint8) if skip is not None else None
        L.check(self._lib, self._lib.chb_arnoldi_step_dev(
            self._plan, batch, n, int(j), int(h.shape[1]), C.c_void_p(V.ctypes.data), C.c_void_p(w.ctypes.data),
            C.c_void_p(sk.ctypes.data) if sk is not None else None, C.c_void_p(h.ctypes.data), None))
        return True

    def _direct_ok(self, a) -> bool:
        """True when the library can address ``a`` directly: CUDA tensors, or NumPy arrays with a build whose ``*_dev`` entry
        points take host pointers (``CHB_Q_HOST_POINTERS``)."""
        return _is_torch(a) or bool(self._q(L.Q_HOST_POINTERS))

    def _ptr_stream(self, like):
        if _is_torch(like):
            import torch
            return (lambda t: C.c_void_p(t.data_ptr()) if t is not None else None), \
                C.c_void_p(torch.cuda.current_stream(like.device).cuda_stream)
        return (lambda t: C.c_void_p(t.ctypes.data) if t is not None else None), None

    def _ws_direct(self, batch, like):
        if _is_torch(like):
            ws = self._workspace(batch, like)
            return C.c_void_p(ws.data_ptr()), ws.numel()
        need = C.c_size_t()
        L.check(self._lib, self._lib.chb_workspace_bytes(self._plan, batch, C.byref(need)))
        if getattr(self, "_ws_np", None) is None or self._ws_np.size < need.value:
            self._ws_np = np.empty(int(need.value), dtype=np.uint8)
        return C.c_void_p(self._ws_np.ctypes.data), self._ws_np.size

    def hess_masked(self, segmpos, dparams, skip, out) -> bool:
        """``segmpos_dparams_to_action_hess`` into ``out`` for the states whose ``skip`` flag (bool ``[batch]``) is clear
        (``chb_segmpos_dparams_to_action_hess_masked_dev``); rows of flagged states may be left untouched.  2-D inputs.
        Returns False when the library cannot address the arrays."""
        if not self._direct_ok(segmpos):
            return False
        b = int(dparams.shape[0])
        if tuple(segmpos.shape) != (b,) + self.segmpos_shape or tuple(dparams.shape) != (b, self.nparams) \
                or tuple(out.shape) != (b, self.nparams) or tuple(skip.shape) != (b,):
            raise ValueError("hess_masked: inconsistent shapes")
        ptr, stream = self._ptr_stream(dparams)
        wsp, wsn = self._ws_direct(b, dparams)
        L.check(self._lib, self._lib.chb_segmpos_dparams_to_action_hess_masked_dev(
            self._plan, b, ptr(segmpos), ptr(dparams), ptr(skip), ptr(out), wsp, wsn, stream))
        return True

    def hess_listed(self, segmpos, dparams, live_list, live_count, out) -> bool:
        """``segmpos_dparams_to_action_hess`` into ``out`` for the states ``live_list[: live_count[0]]`` (int32 arrays in the
        library's memory, e.g. written by ``gmres_step``): ``chb_segmpos_dparams_to_action_hess_listed_dev``."""
        if not self._direct_ok(segmpos):
            return False
        b = int(dparams.shape[0])
        if tuple(segmpos.shape) != (b,) + self.segmpos_shape or tuple(dparams.shape) != (b, self.nparams) \
                or tuple(out.shape) != (b, self.nparams) or int(live_list.shape[0]) < b:
            raise ValueError("hess_listed: inconsistent shapes")
        ptr, stream = self._ptr_stream(dparams)
        wsp, wsn = self._ws_direct(b, dparams)
        L.check(self._lib, self._lib.chb_segmpos_dparams_to_action_hess_listed_dev(
            self._plan, b, ptr(segmpos), ptr(dparams), ptr(live_list), ptr(live_count), ptr(out), wsp, wsn, stream))
        return True

    def row_stats(self, a):
        """(sum of squares, max |.|) of every row of ``a`` ``[batch, n]`` in one pass (``chb_row_stats_dev``), or None when
        the library cannot address the array."""
        if not self._direct_ok(a):
            return None
        b, n = int(a.shape[0]), int(a.shape[1])
        ptr, stream = self._ptr_stream(a)
        if _is_torch(a):
            import torch
            a = a.contiguous()
            ss = torch.empty((b,), dtype=torch.float64, device=a.device)
            mx = torch.empty((b,), dtype=torch.float64, device=a.device)
        else:
            a = np.ascontiguousarray(a, dtype=np.float64)
            ss, mx = np.empty((b,)), np.empty((b,))
        L.check(self._lib, self._lib.chb_row_stats_dev(self._plan, b, n, ptr(a), ptr(ss), ptr(mx), stream))
        return ss, mx

    def gmres_solution(self, V, R, g, kdone, scale: float, out) -> bool:
        """``out[b] = scale * sum_(j < kdone[b]) y_j V[j, b]`` with ``y`` from the back substitution on the state's triangle
        of ``R`` (``chb_gmres_solution_dev``).  Returns False when the library cannot address the arrays."""
        if not self._direct_ok(V):
            return False
        m, batch, n = int(V.shape[0]) - 1, int(V.shape[1]), int(V.shape[2])
        if tuple(R.shape) != (batch, m + 1, m) or tuple(g.shape) != (batch, m + 1) or tuple(kdone.shape) != (batch,) \
                or tuple(out.shape) != (batch, n):
            raise ValueError("gmres_solution: inconsistent shapes")
        ptr, stream = self._ptr_stream(V)
        L.check(self._lib, self._lib.chb_gmres_solution_dev(self._plan, batch, n, m, ptr(V), ptr(R), ptr(g), ptr(kdone),
                                                            float(scale), ptr(out), stream))
        return True

    def gmres_step(self, V, j: int, w, done, R, g, cs, target, kdone, live_list_next=None, live_count_next=None) -> bool:
        """One whole inner iteration of the batched GMRES (``chb_gmres_step_dev``): Arnoldi step + Givens rotations + R / g /
        kdone / done updates, all in place, one launch.  Shapes: ``V [m+1, batch, n]``, ``w [batch, n]``, ``done [batch]``
        bool, ``R [batch, m+1, m]``, ``g [batch, m+1]``, ``cs [batch, m, 2]``, ``target [batch]``, ``kdone [batch]`` int64.
        ``live_list_next`` (int32 ``[batch]``) / ``live_count_next`` (int32 ``[1]``, zero on entry): the states still running
        after this step are appended, for ``hess_listed`` of the next step.
        Returns False when the library cannot address the arrays (see ``arnoldi_step``)."""
        torch_in = _is_torch(V)
        if not self._direct_ok(V):
            return False
        batch, n = int(w.shape[0]), int(w.shape[1])
        m = int(R.shape[2])
        if m > 62:
            return False
        want = {"V": (m + 1, batch, n), "R": (batch, m + 1, m), "g": (batch, m + 1), "cs": (batch, m, 2), "target": (batch,),
                "kdone": (batch,), "done": (batch,)}
        got = {"V": V, "R": R, "g": g, "cs": cs, "target": target, "kdone": kdone, "done": done}
        for k, shp in want.items():
            if tuple(got[k].shape) != shp:
                raise ValueError(f"gmres_step: {k} has shape {tuple(got[k].shape)}, expected {shp}")
        if torch_in:
            import torch
            for k, t in got.items():
                dt = torch.bool if k == "done" else torch.int64 if k == "kdone" else torch.float64
                if t.dtype != dt or not t.is_cuda or not t.is_contiguous():
                    raise ValueError(f"gmres_step: {k} must be a contiguous {dt} CUDA tensor")
            if w.dtype != torch.float64 or not w.is_contiguous():
                raise ValueError("gmres_step: w must be a contiguous float64 CUDA tensor")
            ptr = lambda t: C.c_void_p(t.data_ptr())
            stream = C.c_void_p(torch.cuda.current_stream(V.device).cuda_stream)
        else:
            for k, t in got.items():
                dt = np.bool_ if k == "done" else np.int64 if k == "kdone" else np.float64
                if t.dtype != dt or not t.flags.c_contiguous:
                    raise ValueError(f"gmres_step: {k} must be a C-contiguous {dt} array")
            if w.dtype != np.float64 or not w.flags.c_contiguous:
                raise ValueError("gmres_step: w must be a C-contiguous float64 array")
            ptr = lambda t: C.c_void_p(t.ctypes.data)
            stream = None
        L.check(self._lib, self._lib.chb_gmres_step_dev(self._plan, batch, n, int(j), m, ptr(V), ptr(w), ptr(done), ptr(R),
                                                        ptr(g), ptr(cs), ptr(target), ptr(kdone),
                                                        ptr(live_list_next) if live_list_next is not None else None,
                                                        ptr(live_count_next) if live_count_next is not None else None, stream))
        return True

    def newton_krylov(self, x, f_tol: float, maxiter: int, restart: int = 20, smin: float = 1e-2, line_search: Optional[str] = "armijo",
                      tol_norm: str = "max", track_best: bool = False):
        """The whole batched Newton-Krylov solve inside the library (``chb_newton_krylov_dev``): ``x`` ``[batch, nparams]`` is
        updated IN PLACE (torch CUDA tensor, or a NumPy array with the emulation build).  Returns a dict with ``status``,
        ``nit`` (int32 arrays of the carrier), ``fnorm_max``, ``fnorm_2``, ``best_x`` / ``best_fnorm`` (when ``track_best``)
        and the counters ``n_grad_evals``, ``n_hess_evals``, ``launches``, ``iterations``; None when the library cannot
        address ``x``."""
        if not self._direct_ok(x):
            return None
        ls = {None: 0, "armijo": 1, "wolfe": 2}
        if line_search not in ls:
            raise ValueError(f"line_search must be None, 'armijo' or 'wolfe' (nonlin.py:129), got {line_search!r}")
        if tol_norm not in ("max", "l2"):
            raise ValueError("tol_norm must be 'max' or 'l2'")
        b, n = int(x.shape[0]), int(x.shape[1])
        if n != self.nparams:
            raise ValueError(f"x: expected trailing shape ({self.nparams},), got {tuple(x.shape)}")
        opts = L.chb_newton_opts(float(f_tol), float(smin), int(maxiter), int(restart), ls[line_search], 1 if tol_norm == "l2" else 0)
        need = C.c_size_t()
        L.check(self._lib, self._lib.chb_newton_workspace_bytes(self._plan, b, int(restart), C.byref(need)))
        counters = (C.c_int64 * 4)()
        if _is_torch(x):
            import torch
            if x.dtype != torch.float64 or not x.is_cuda or not x.is_contiguous():
                raise ValueError("newton_krylov: x must be a contiguous float64 CUDA tensor")
            mk = lambda shape, dt: torch.empty(shape, dtype=dt, device=x.device)  # noqa: E731
            i32, f64 = torch.int32, torch.float64
            key = ("nk_ws", str(x.device))
            ws = getattr(self, "_nk_ws", None)
            if ws is None or ws.numel() < need.value or ws.device != x.device:
                self._nk_ws = None
                ws = self._nk_ws = torch.empty(int(need.value), dtype=torch.uint8, device=x.device)
            vp = lambda t: C.c_void_p(t.data_ptr()) if t is not None else None  # noqa: E731
            stream = C.c_void_p(torch.cuda.current_stream(x.device).cuda_stream)
            wsp, wsn = C.c_void_p(ws.data_ptr()), ws.numel()
        else:
            if x.dtype != np.float64 or not x.flags.c_contiguous:
                raise ValueError("newton_krylov: x must be a C-contiguous float64 array")
            mk = lambda shape, dt: np.empty(shape, dtype=dt)  # noqa: E731
            i32, f64 = np.int32, np.float64
            ws = np.empty(int(need.value), dtype=np.uint8)
            vp = lambda t: C.c_void_p(t.ctypes.data) if t is not None else None  # noqa: E731
            stream = None
            wsp, wsn = C.c_void_p(ws.ctypes.data), ws.size
        status, nit = mk((b,), i32), mk((b,), i32)
        fmax, f2 = mk((b,), f64), mk((b,), f64)
        best_x = mk((b, n), f64) if track_best else None
        best_f = mk((b,), f64) if track_best else None
        L.check(self._lib, self._lib.chb_newton_krylov_dev(self._plan, b, vp(x), C.byref(opts), vp(status), vp(nit), vp(fmax), vp(f2),
                                                            vp(best_x), vp(best_f), counters, wsp, wsn, stream))
        return {"status": status, "nit": nit, "fnorm_max": fmax, "fnorm_2": f2, "best_x": best_x, "best_fnorm": best_f,
                "n_grad_evals": int(counters[0]), "n_hess_evals": int(counters[1]), "launches": int(counters[2]),
                "iterations": int(counters[3])}

    # ------------------------------------------------------------------ post-solve reductions (find.py:285-308, 376-380)
    def ComputeCenterOfMass(self, segmpos):
        """_NBodySyst.pyx:3968-4009, batched (``chb_center_of_mass_dev``, one CTA per state): mass-weighted mean position over
        all bodies and the whole period, from the stored segments and the symmetry that carries each body onto them.
        ``[..., nsegm, segm_store, geodim] -> [..., geodim]``."""
        self._need_body_tables()
        x, b, lead = self._prep(segmpos, self.segmpos_shape, "segmpos")
        if b == 0:
            return (x.new_empty if _is_torch(x) else np.empty)(lead + (self.geodim,))
        out, = self._arr_call("chb_center_of_mass_dev", [x], [(lead + (self.geodim,), np.float64)], lambda i, o: (b, i[0], o[0]))
        return out

    def DetectEscape(self, segmpos, fac: float = 2.):
        """_NBodySyst.pyx:2125-2192, batched (``chb_detect_escape_dev``, one CTA per state): per body the box spanned by the
        rotated min / max corners of its segments, an edge between two bodies whose box centres are closer than ``fac`` times
        the sum of the box diagonals; escaped = that graph is not connected.  Returns a bool (one state), a NumPy bool array
        (NumPy input) or a bool tensor (device input) over the leading dimensions."""
        self._need_body_tables()
        x, b, lead = self._prep(segmpos, self.segmpos_shape, "segmpos")
        if b == 0:
            return np.zeros(lead, dtype=bool)
        out, = self._arr_call("chb_detect_escape_dev", [x], [(lead if lead else (1,), np.int32)],
                              lambda i, o: (b, i[0], float(fac), o[0]))
        esc = out != 0
        if lead == ():
            return bool(esc.reshape(-1)[0])
        return esc

    def segm_to_path_stats(self, segmpos, segmvel):
        """_NBodySyst.pyx:3903-3943 (cdef :7420-7529), batched (``chb_path_stats_dev``, one CTA per state): length of every loop
        (trapezoid rule on ``|v|`` when the segments store their end point, plain sum otherwise) and the minimum distance of
        every binary interaction over the stored samples.  ``segmvel`` has the shape of ``segmpos`` (``params_to_segmvel``).
        Returns ``(loop_len [..., nloop], bin_dx_min [..., nbin])``."""
        x, b, lead = self._prep(segmpos, self.segmpos_shape, "segmpos")
        v, bv, _ = self._prep(segmvel, self.segmpos_shape, "segmvel")
        if b != bv:
            raise ValueError("segmpos and segmvel hold different numbers of states")
        nbin = self.tables.nbin
        if b == 0:
            mk = x.new_empty if _is_torch(x) else np.empty
            return mk(lead + (self.nloop,)), mk(lead + (nbin,))
        ll, bm = self._arr_call("chb_path_stats_dev", [x, v], [(lead + (self.nloop,), np.float64), (lead + (nbin,), np.float64)],
                                lambda i, o: (b, i[0], i[1], o[0], o[1]))
        return ll, bm

    def _need_body_tables(self):
        if self.tables.bodysegm is None:
            raise ValueError("these tables carry no per-body symmetries (bodysegm / body_spacerot / body_timerev): "
                             "re-extract them with extract_tables from a live NBodySyst")

    @property
    def LawIsHomo(self) -> bool:
        return True  # gravity_pot and power_law_pot are homogeneous (Detect_homo_inter_law, _NBodySyst.pyx:1692-1718)

    @property
    def Homo_exp(self) -> float:
        """``xsq V'(xsq) / V(xsq)`` (Detect_homo_inter_law, _NBodySyst.pyx:1692-1718): -1/2 for gravity, n/2 for the power law."""
        if self.tables.inter_law == 0:
            return -0.5
        return float(self.tables.inter_law_params[2]) + 2.0

    def TestHashSame(self, Hash_a, Hash_b, rtol: float = 1e-5, detect_multiples: bool = True) -> bool:
        """_NBodySyst.pyx:2036-2067.  For a homogeneous law (gravity / power law: Homo_exp = n / 2) two orbits that differ by a
        rescaling have hashes whose ratios follow a power law; otherwise the hashes must agree entry by entry."""
        Ha, Hb = np.asarray(Hash_a, dtype=np.float64), np.asarray(Hash_b, dtype=np.float64)
        n = self.Hash_exp.shape[0]
        test = np.zeros(n)
        if detect_multiples:
            homo = self.Homo_exp
            ref = (Ha[0] / Hb[0]) ** ((homo - 1) / (2 * self.Hash_exp[0]))
            for i in range(1, n):
                test[i] = (Ha[i] / Hb[i]) ** ((homo - 1) / (2 * self.Hash_exp[i])) - ref
        else:
            test = (Hb - Ha) / (Hb + Ha)
        return bool(np.linalg.norm(test, np.inf) < rtol)

    def segmpos_to_pot_nrg_grad(self, segmpos, globalmul: float = 1.0):
        """cdef segm_pos_to_pot_nrg_grad, _NBodySyst.pyx:6472-6533 (into a zeroed buffer)."""
        if _is_torch(segmpos):
            import torch
            x, b, lead = self._tprep(segmpos, self.segmpos_shape, "segmpos")
            out = torch.empty_like(x)
            self._dev("chb_segmpos_to_pot_nrg_grad_dev", b, [x, float(globalmul), out], x)
            return out
        return self._host(L.OP_POT_GRAD, segmpos, self.segmpos_shape, out0_shape=self.segmpos_shape,
                          scalar=globalmul)[0]

    def segmpos_to_pot_nrg_hess(self, segmpos, dsegmpos, globalmul: float = 1.0):
        """cdef segm_pos_to_pot_nrg_hess, _NBodySyst.pyx:6899-6962."""
        if _is_torch(segmpos):
            import torch
            x, b, lead = self._tprep(segmpos, self.segmpos_shape, "segmpos")
            dx, b2, _ = self._tprep(dsegmpos, self.segmpos_shape, "dsegmpos")
            if b != b2:
                raise ValueError("batch sizes differ")
            out = torch.empty_like(x)
            self._dev("chb_segmpos_to_pot_nrg_hess_dev", b, [x, dx, float(globalmul), out], x)
            return out
        return self._host(L.OP_POT_HESS, segmpos, self.segmpos_shape, dsegmpos, self.segmpos_shape,
                          out0_shape=self.segmpos_shape, scalar=globalmul)[0]

    def params_to_pot_nrg(self, params_mom_buf):
        """_NBodySyst.pyx:2563-2600."""
        return self.segmpos_to_pot_nrg(self.params_to_segmpos(params_mom_buf))

    def params_to_pot_nrg_grad(self, params_mom_buf):
        """_NBodySyst.pyx:2603-2661."""
        return self.segmpos_to_params_T(self.segmpos_to_pot_nrg_grad(self.params_to_segmpos(params_mom_buf), 1.0))

    def params_to_pot_nrg_hess(self, params_mom_buf, dparams_mom_buf):
        """_NBodySyst.pyx:2664-2747."""
        seg = self.params_to_segmpos(params_mom_buf)
        dseg = self.params_to_segmpos(dparams_mom_buf)
        return self.segmpos_to_params_T(self.segmpos_to_pot_nrg_hess(seg, dseg, 1.0))

    # ------------------------------------------------------------------ action chain
    def params_to_action(self, params_mom_buf):
        """_NBodySyst.pyx:2750-2791."""
        if _is_torch(params_mom_buf):
            import torch
            x, b, lead = self._tprep(params_mom_buf, (self.nparams,), "params")
            out = torch.empty(lead, dtype=torch.float64, device=x.device)
            self._dev("chb_params_to_action_dev", b, [x, out], x)
            return out
        a = np.asarray(params_mom_buf)
        o, _ = self._host(L.OP_ACTION, a, (self.nparams,), out0_shape=())
        return self._scalar(o, a.shape[:-1])

    def params_to_action_grad(self, params_mom_buf):
        """_NBodySyst.pyx:2794-2873."""
        return self.params_to_action_and_grad(params_mom_buf, want_action=False)[1]

    def params_to_action_and_grad(self, params_mom_buf, want_action: bool = True, want_segmpos: bool = False,
                                  out_action=None, out_grad=None):
        """One fused evaluation of ``params_to_action`` + ``params_to_action_grad`` (unit U1 of SURVEY.md 8(d)).
        Returns (action or None, grad[, segmpos]).  Host path: ``out_action`` / ``out_grad`` may be preallocated
        (e.g. ``pinned_empty``) arrays that receive the results."""
        if _is_torch(params_mom_buf):
            import torch
            x, b, lead = self._tprep(params_mom_buf, (self.nparams,), "params")
            act = torch.empty(lead, dtype=torch.float64, device=x.device) if want_action else None
            grad = torch.empty_like(x)
            seg = torch.empty(lead + self.segmpos_shape, dtype=torch.float64, device=x.device) if want_segmpos else None
            self._dev("chb_params_to_action_grad_dev", b, [x, act, grad, seg], x)
            return (act, grad, seg) if want_segmpos else (act, grad)
        a = np.asarray(params_mom_buf)
        if want_segmpos:
            raise ValueError("want_segmpos is only available on the device (torch) path")
        o0, o1 = self._host(L.OP_ACTION_GRAD, a, (self.nparams,), out0_shape=(), out1_shape=(self.nparams,),
                            want_out0=want_action, out0=out_action, out1=out_grad)
        return (self._scalar(o0, a.shape[:-1]) if want_action else None), o1

    def params_to_action_hess(self, params_mom_buf, dparams_mom_buf):
        """_NBodySyst.pyx:2876-2965."""
        if _is_torch(params_mom_buf):
            import torch
            x, b, lead = self._tprep(params_mom_buf, (self.nparams,), "params")
            dx, b2, _ = self._tprep(dparams_mom_buf, (self.nparams,), "dparams")
            if b != b2:
                raise ValueError("batch sizes differ")
            out = torch.empty_like(x)
            self._dev("chb_params_to_action_hess_dev", b, [x, dx, out], x)
            return out
        return self._host(L.OP_ACTION_HESS, params_mom_buf, (self.nparams,), dparams_mom_buf, (self.nparams,),
                          out0_shape=(self.nparams,))[0]

    NOALLOC_SCRATCH_BYTES = 32 << 20  # CHB_NOALLOC_SCRATCH_BYTES of include/choreo_b200.h

    def params_to_action_hess_noalloc(self, params_mom_buf, dparams_mom_buf, out) -> None:
        """``out[...] = params_to_action_hess(params, dparams)`` for callers that cannot allocate device memory (CUDA graph capture,
        serving loops): no counterpart in the reference.  All three arguments are
        contiguous float64 CUDA tensors of shape ``[batch, nparams]`` (or ``[nparams]``) on the plan's device; nothing is allocated
        here or in the library (``chb_params_to_action_hess_noalloc_dev``: the scratch is a 32 MiB device array built into the
        library, shared by every call on the device; a batch that needs more raises ``ChoreoB200Error`` CHB_ERR_WORKSPACE;
        calls that overlap on the GPU race on it)."""
        import torch
        for name, t in (("params", params_mom_buf), ("dparams", dparams_mom_buf), ("out", out)):
            if not (_is_torch(t) and t.is_cuda and t.dtype == torch.float64 and t.is_contiguous()):
                raise ValueError(f"params_to_action_hess_noalloc: {name} must be a contiguous float64 CUDA tensor")
            if t.shape[-1] != self.nparams or t.shape != params_mom_buf.shape:
                raise ValueError(f"params_to_action_hess_noalloc: {name} has shape {tuple(t.shape)}, expected [batch, {self.nparams}]")
        b = 1 if params_mom_buf.dim() == 1 else int(params_mom_buf.numel() // self.nparams)
        stream = torch.cuda.current_stream(params_mom_buf.device).cuda_stream
        L.check(self._lib, self._lib.chb_params_to_action_hess_noalloc_dev(
            self._plan, b, C.c_void_p(params_mom_buf.data_ptr()), C.c_void_p(dparams_mom_buf.data_ptr()), C.c_void_p(out.data_ptr()),
            C.c_void_p(stream)))

    def segmpos_params_to_action(self, segmpos, params_mom_buf):
        """_NBodySyst.pyx:2968-2991."""
        if _is_torch(segmpos):
            import torch
            s, b, lead = self._tprep(segmpos, self.segmpos_shape, "segmpos")
            x, b2, _ = self._tprep(params_mom_buf, (self.nparams,), "params")
            if b != b2:
                raise ValueError("batch sizes differ")
            out = torch.empty(lead, dtype=torch.float64, device=x.device)
            self._dev("chb_segmpos_params_to_action_dev", b, [s, x, out], x)
            return out
        a = np.asarray(params_mom_buf)
        o, _ = self._host(L.OP_SEGMPOS_ACTION, segmpos, self.segmpos_shape, a, (self.nparams,), out0_shape=())
        return self._scalar(o, a.shape[:-1])

    def segmpos_params_to_action_grad(self, segmpos, params_mom_buf):
        """_NBodySyst.pyx:2994-3041."""
        if _is_torch(segmpos):
            import torch
            s, b, lead = self._tprep(segmpos, self.segmpos_shape, "segmpos")
            x, b2, _ = self._tprep(params_mom_buf, (self.nparams,), "params")
            if b != b2:
                raise ValueError("batch sizes differ")
            out = torch.empty_like(x)
            self._dev("chb_segmpos_params_to_action_grad_dev", b, [s, x, None, out], x)
            return out
        return self._host(L.OP_SEGMPOS_ACTION_GRAD, segmpos, self.segmpos_shape, params_mom_buf, (self.nparams,),
                          out0_shape=(), out1_shape=(self.nparams,), want_out0=False)[1]

    def segmpos_dparams_to_action_hess(self, segmpos, dparams_mom_buf):
        """_NBodySyst.pyx:3044-3110 (the Krylov matvec)."""
        if _is_torch(segmpos):
            import torch
            s, b, lead = self._tprep(segmpos, self.segmpos_shape, "segmpos")
            x, b2, _ = self._tprep(dparams_mom_buf, (self.nparams,), "dparams")
            if b != b2:
                raise ValueError("batch sizes differ")
            out = torch.empty_like(x)
            self._dev("chb_segmpos_dparams_to_action_hess_dev", b, [s, x, out], x)
            return out
        return self._host(L.OP_SEGMPOS_ACTION_HESS, segmpos, self.segmpos_shape, dparams_mom_buf, (self.nparams,),
                          out0_shape=(self.nparams,))[0]

    # batched aliases: every method above already accepts leading batch dimensions
    params_to_action_batched = params_to_action
    params_to_action_grad_batched = params_to_action_grad
    params_to_action_hess_batched = params_to_action_hess
    params_to_segmpos_batched = params_to_segmpos
    segmpos_dparams_to_action_hess_batched = segmpos_dparams_to_action_hess

    # ------------------------------------------------------------------ solver glue
    def GetKrylovJacobian(self, Use_exact_Jacobian: bool = True, jac_options_kw: Optional[dict] = None):
        """_NBodySyst.pyx:2004-2032: scipy KrylovJacobian whose matvec is the exact Hessian-vector product with
        cached segment positions."""
        try:
            from scipy.optimize._nonlin import KrylovJacobian
        except Exception:  # scipy < 1.14
            from scipy.optimize.nonlin import KrylovJacobian
        jacobian = KrylovJacobian(**(jac_options_kw or {}))
        if Use_exact_Jacobian:
            def matvec(jself, v):
                return jself.NBS.segmpos_dparams_to_action_hess(jself.segmpos, v)

            def update(jself, x, f):
                jself.segmpos = jself.NBS.params_to_segmpos(x)
                KrylovJacobian.update(jself, x, f)

            def setup(jself, x, f, func):
                jself.segmpos = jself.NBS.params_to_segmpos(x)
                KrylovJacobian.setup(jself, x, f, func)

            jacobian.matvec = types.MethodType(matvec, jacobian)
            jacobian.rmatvec = types.MethodType(matvec, jacobian)
            jacobian.update = types.MethodType(update, jacobian)
            jacobian.setup = types.MethodType(setup, jacobian)
        jacobian.NBS = self
        return jacobian
