"""Symmetry tables ("the plan input") consumed by the B200 action hot path.

The reference builds these tables once per system in ``NBodySyst.__init__``
(choreo/cython/_NBodySyst.pyx:742-896, DetectLoops ... Compute_n_sub_fft) and its numeric
kernels only ever read them.  Symmetry setup itself is out of scope here (SURVEY.md section 2,
rows 2/5/6): we *extract* the tables from a live reference ``NBodySyst`` (``extract_tables``) or
load them from an ``.npz`` written by ``PlanTables.save`` and hand them to the C ABI
(``chb_plan_create``, include/choreo_b200.h).

Only public attributes of the reference object are touched, following the pitfalls listed in
SURVEY.md section 8(b): position rotations are re-derived from ``intersegm_to_all`` /
``gensegm_to_all`` (``NBS.InterSpaceRot`` returns the *velocity* variant,
_NBodySyst.pyx:401-419) and ``NBS.gensegm_loop_start`` is never touched (it recurses, :362-364).
"""
from __future__ import annotations

import dataclasses
import io
import os
from typing import Optional, Any, Dict, List, Tuple

import numpy as np

GENERAL_SYM = 0  # _NBodySyst.pyx:94
RFFT = 1  # _NBodySyst.pyx:95

LAW_GRAVITY = 0  # gravity_pot, _NBodySyst.pyx:111-123
LAW_POWER = 1  # power_law_pot, _NBodySyst.pyx:125-133

_INT = np.int64
_I32 = np.int32
_F64 = np.float64
_C128 = np.complex128


@dataclasses.dataclass
class LevelShapes:
    """Everything that depends on ``nint`` (reference: ``nint`` setter, _NBodySyst.pyx:540-590)."""

    nint: int
    segm_size: int
    segm_store: int
    npr: np.ndarray  # [nloop]   params_shapes[il,0]
    ninter: np.ndarray  # [nloop]   pos_slice_shapes[il,0]
    nmem: np.ndarray  # [nloop]   npr*nnz*nppl  (pos-params incl. removed zero-frequency entries)
    pos_off: np.ndarray  # [nloop+1] params_shifts (in pos-param space)
    mom_off: np.ndarray  # [nloop+1] offsets in the user-visible parameter vector
    slice_off: np.ndarray  # [nloop+1] pos_slice_shifts (in doubles)
    nparams: int
    nparams_incl_o: int


@dataclasses.dataclass
class PlanTables:
    # scalars
    geodim: int
    nbody: int
    nloop: int
    nsegm: int
    nbin: int
    nint_min: int
    greater_nstore: bool
    inter_law: int
    inter_law_params: np.ndarray  # [4]  (-nu, -nu(nu-1), nu-2, alpha), _NBodySyst.pyx:1623-1629
    # per loop
    loopnb: np.ndarray
    loopmass: np.ndarray
    ncoeff_min: np.ndarray
    n_sub_fft: np.ndarray
    shortcut: np.ndarray
    nnz: np.ndarray
    nppl: np.ndarray
    nnz_k: np.ndarray
    nnz_k_shifts: np.ndarray
    co_in: np.ndarray
    co_in_shifts: np.ndarray
    params_basis: np.ndarray  # complex128 flat, per loop C-order (geodim, nnz, nppl)
    params_basis_shifts: np.ndarray
    alg_iint: np.ndarray
    alg_timerev: np.ndarray
    alg_spacerot: np.ndarray  # [nloop, G, G]
    # per body
    bodyloop: np.ndarray
    # per segment
    gensegm_to_body: np.ndarray
    gensegm_to_iintrel: np.ndarray
    inter_timerev: np.ndarray
    inter_spacerot: np.ndarray  # [nsegm, G, G]  (position variant)
    inter_spacerot_isid: np.ndarray
    # per binary interaction
    bin_source: np.ndarray
    bin_target: np.ndarray
    bin_spacerot: np.ndarray  # [nbin, G, G]
    bin_spacerot_isid: np.ndarray
    bin_prodchargesum: np.ndarray
    # per (body, interval of the period): which stored segment the body runs on and the symmetry that maps it there
    # (``_bodysegm`` and ``intersegm_to_all[ib][iint]``) -- only the post-solve reductions (centre of mass, escape test) use
    # them; optional so that older table files keep loading
    bodysegm: Optional[np.ndarray] = None       # [nbody, nint_min]
    body_spacerot: Optional[np.ndarray] = None  # [nbody, nint_min, G, G]
    body_timerev: Optional[np.ndarray] = None   # [nbody, nint_min]

    # ------------------------------------------------------------------ helpers
    def __post_init__(self) -> None:
        c = np.ascontiguousarray
        self.inter_law_params = c(self.inter_law_params, dtype=_F64)
        for name in ("loopnb", "ncoeff_min", "n_sub_fft", "nnz", "nppl", "nnz_k", "nnz_k_shifts",
                     "co_in_shifts", "params_basis_shifts", "alg_iint", "alg_timerev", "bodyloop",
                     "gensegm_to_body", "gensegm_to_iintrel", "inter_timerev", "bin_source", "bin_target"):
            setattr(self, name, c(getattr(self, name), dtype=_INT))
        for name in ("shortcut", "co_in", "inter_spacerot_isid", "bin_spacerot_isid"):
            setattr(self, name, c(getattr(self, name), dtype=_I32))
        for name in ("loopmass", "alg_spacerot", "inter_spacerot", "bin_spacerot", "bin_prodchargesum"):
            setattr(self, name, c(getattr(self, name), dtype=_F64))
        self.params_basis = c(self.params_basis, dtype=_C128)
        G = self.geodim
        self.alg_spacerot = self.alg_spacerot.reshape(self.nloop, G, G)
        self.inter_spacerot = self.inter_spacerot.reshape(self.nsegm, G, G)
        self.bin_spacerot = self.bin_spacerot.reshape(self.nbin, G, G)
        if self.bodysegm is not None:
            self.bodysegm = c(self.bodysegm, dtype=_INT).reshape(self.nbody, self.nint_min)
            self.body_timerev = c(self.body_timerev, dtype=_INT).reshape(self.nbody, self.nint_min)
            self.body_spacerot = c(self.body_spacerot, dtype=_F64).reshape(self.nbody, self.nint_min, G, G)

    def segm_loop(self) -> np.ndarray:
        """Loop index of every generating segment (BodyLoop[gensegm_to_body], _NBodySyst.pyx:5774-5776)."""
        return self.bodyloop[self.gensegm_to_body]

    def loop_nnz_k(self, il: int) -> np.ndarray:
        return self.nnz_k[self.nnz_k_shifts[il]:self.nnz_k_shifts[il + 1]]

    def loop_co_in(self, il: int) -> np.ndarray:
        return self.co_in[self.co_in_shifts[il]:self.co_in_shifts[il + 1]]

    def loop_basis(self, il: int) -> np.ndarray:
        b = self.params_basis[self.params_basis_shifts[il]:self.params_basis_shifts[il + 1]]
        return b.reshape(self.geodim, int(self.nnz[il]), int(self.nppl[il]))

    def loop_has_zero_freq(self, il: int) -> bool:
        return bool(self.nnz[il] > 0 and self.loop_nnz_k(il)[0] == 0)

    def ncor_nco_in(self, il: int) -> Tuple[int, int]:
        """(kept, removed) zero-frequency parameters of loop il (_NBodySyst.pyx:864-871)."""
        ci = self.loop_co_in(il)
        ncor = int(np.count_nonzero(ci))
        return ncor, int(ci.shape[0] - ncor)

    def check_nint(self, nint: int) -> None:
        if nint <= 0 or (nint % (2 * self.nint_min)) != 0:
            # same message as the reference nint setter, _NBodySyst.pyx:545-546
            raise ValueError(f"Provided nint {nint} should be divisible by {2 * self.nint_min}")

    def level(self, nint: int) -> LevelShapes:
        self.check_nint(nint)
        segm_size = nint // self.nint_min
        segm_store = segm_size + (1 if self.greater_nstore else 0)
        npr = np.zeros(self.nloop, dtype=_INT)
        ninter = np.zeros(self.nloop, dtype=_INT)
        nmem = np.zeros(self.nloop, dtype=_INT)
        pos_off = np.zeros(self.nloop + 1, dtype=_INT)
        mom_off = np.zeros(self.nloop + 1, dtype=_INT)
        slice_off = np.zeros(self.nloop + 1, dtype=_INT)
        for il in range(self.nloop):
            ncm = int(self.ncoeff_min[il])
            if nint % (2 * ncm) != 0:
                raise ValueError(f"nint {nint} is not divisible by 2*ncoeff_min_loop[{il}] = {2 * ncm}")
            npr[il] = nint // (2 * ncm)
            if self.n_sub_fft[il] == 2:
                ninter[il] = npr[il] + 1
            elif self.n_sub_fft[il] == 1:
                ninter[il] = 2 * npr[il] + (1 if self.greater_nstore else 0)
            else:
                raise ValueError(f"Impossible value for n_sub_fft[il]: {self.n_sub_fft[il]}. "
                                 "Allowed values are either 1 or 2.")
            nmem[il] = npr[il] * self.nnz[il] * self.nppl[il]
            pos_off[il + 1] = pos_off[il] + nmem[il]
            mom_off[il + 1] = mom_off[il] + nmem[il] - self.ncor_nco_in(il)[1]
            slice_off[il + 1] = slice_off[il] + ninter[il] * self.geodim
        return LevelShapes(nint=nint, segm_size=segm_size, segm_store=segm_store, npr=npr, ninter=ninter,
                           nmem=nmem, pos_off=pos_off, mom_off=mom_off, slice_off=slice_off,
                           nparams=int(mom_off[-1]), nparams_incl_o=int(2 * pos_off[-1]))

    # ------------------------------------------------------------------ (de)serialisation
    _SCALARS = ("geodim", "nbody", "nloop", "nsegm", "nbin", "nint_min", "greater_nstore", "inter_law")

    def params_resize(self, params, nint_in: int, nint_out: int):
        """Batched restatement of ``NBodySyst.params_resize`` (_NBodySyst.pyx:2413-2465): per loop the leading (lowest
        frequency) entries are copied, the new high-frequency entries are zero (refinement) or the old ones dropped
        (coarsening).  ``params``: ``[..., nparams(nint_in)]`` NumPy array or torch tensor; returns ``[..., nparams(nint_out)]``."""
        lin, lout = self.level(int(nint_in)), self.level(int(nint_out))
        if params.shape[-1] != lin.nparams:
            raise ValueError(f"params_resize: last dimension {params.shape[-1]}, expected nparams = {lin.nparams}")
        shape = tuple(params.shape[:-1]) + (lout.nparams,)
        if type(params).__module__.startswith("torch"):
            import torch
            out = torch.zeros(shape, dtype=params.dtype, device=params.device)
        else:
            out = np.zeros(shape, dtype=np.asarray(params).dtype)
        for il in range(self.nloop):
            n = int(min(lin.mom_off[il + 1] - lin.mom_off[il], lout.mom_off[il + 1] - lout.mom_off[il]))
            out[..., int(lout.mom_off[il]):int(lout.mom_off[il]) + n] = params[..., int(lin.mom_off[il]):int(lin.mom_off[il]) + n]
        return out

    def make_params_bounds(self, nint: int, coeff_ampl_o: float = 1e-1, k_infl: int = 1, k_max: int = 200,
                           coeff_ampl_min: float = 1e-16) -> Tuple[np.ndarray, np.ndarray]:
        """``NBodySyst.Make_params_bounds`` (_NBodySyst.pyx:2301-2327): amplitude envelope of ``Make_Init_bounds_coeffs``
        (:4447-4486; flat up to harmonic ``k_infl``, then exponential decay reaching ``coeff_ampl_min`` at ``k_max``) mapped to
        the "mom" scaling by ``changevar_mom_pos_inv`` (:4662-4743).  Returns (x_min, x_max), each ``[nparams]``; the
        reference draws initial guesses as ``x_min + (x_max - x_min) * U[0,1)`` (find.py:105-108,158)."""
        lvl = self.level(nint)
        slope = np.log(abs(coeff_ampl_o / coeff_ampl_min)) / (k_max - k_infl)
        out = np.zeros(int(lvl.mom_off[-1]), dtype=np.float64)
        for il in range(self.nloop):
            ncm, nnz, nppl = int(self.ncoeff_min[il]), int(self.nnz[il]), int(self.nppl[il])
            if nnz == 0:
                continue
            k = (np.arange(int(lvl.npr[il]))[:, None] * ncm + self.loop_nnz_k(il)[None, :]).astype(np.float64)
            ampl = np.where(k <= k_infl, coeff_ampl_o, coeff_ampl_o * np.exp(-slope * (k - k_infl)))
            loopmul = np.sqrt(float(self.loopnb[il]) * float(self.loopmass[il]) * 4.0 * np.pi ** 2)
            mom = ampl * np.where(k > 0, loopmul * k, loopmul)
            full = np.repeat(mom[:, :, None], nppl, axis=2).reshape(-1)
            if self.loop_has_zero_freq(il):  # the co_in == False entries of (ipr = 0, ik = 0) are not parameters
                keep = np.ones(full.shape[0], dtype=bool)
                keep[:nppl] = self.loop_co_in(il).astype(bool)
                full = full[keep]
            out[int(lvl.mom_off[il]):int(lvl.mom_off[il + 1])] = full
        return -out, out

    def to_dict(self) -> Dict[str, np.ndarray]:
        out: Dict[str, np.ndarray] = {}
        for f in dataclasses.fields(self):
            v = getattr(self, f.name)
            if v is None:
                continue
            out[f.name] = np.asarray(int(v)) if f.name in self._SCALARS else np.asarray(v)
        return out

    @classmethod
    def from_dict(cls, d: Dict[str, Any], prefix: str = "") -> "PlanTables":
        kw = {}
        for f in dataclasses.fields(cls):
            if f.default is None and (prefix + f.name) not in d:
                continue
            v = d[prefix + f.name]
            if f.name in cls._SCALARS:
                v = int(np.asarray(v))
                if f.name == "greater_nstore":
                    v = bool(v)
            kw[f.name] = v
        return cls(**kw)

    def save(self, path: str) -> None:
        np.savez_compressed(path, **self.to_dict())

    @classmethod
    def load(cls, path: str) -> "PlanTables":
        with np.load(path) as z:
            return cls.from_dict({k: z[k] for k in z.files})

    def with_forced_general_sym(self) -> "PlanTables":
        """Equivalent of ``NBS.ForceGeneralSym = True`` (_NBodySyst.pyx:708-724)."""
        return dataclasses.replace(self, shortcut=np.zeros(self.nloop, dtype=_I32))


def _law_from_reference(NBS) -> Tuple[int, np.ndarray]:
    name = NBS.inter_law_str
    if name == "gravity_pot":
        return LAW_GRAVITY, np.array([0.5, -0.75, -2.5, 1.0])
    if name == "power_law_pot":
        d = NBS.inter_law_param_dict
        nu = d["n"] / 2  # argument is xsq, _NBodySyst.pyx:1623
        return LAW_POWER, np.array([-nu, -nu * (nu - 1), nu - 2, d["alpha"]], dtype=_F64)
    raise ValueError(
        f"Unsupported inter_law {name!r}: the B200 path implements 'gravity_pot' and 'power_law_pot' only "
        "(custom C / numba callbacks cannot run on the device).")


def extract_tables(NBS) -> PlanTables:
    """Read the hot-path tables out of a live reference ``choreo.NBodySyst``."""
    G = int(NBS.geodim)
    nloop = int(NBS.nloop)
    nsegm = int(NBS.nsegm)
    nint_min = int(NBS.nint_min)

    nnz_k_list = [np.asarray(NBS.nnz_k(il), dtype=_INT).reshape(-1) for il in range(nloop)]
    co_in_list = [np.asarray(NBS.co_in(il)).astype(_I32).reshape(-1) for il in range(nloop)]
    basis_list = [np.asarray(NBS.params_basis_pos(il), dtype=_C128) for il in range(nloop)]

    def shifts(lst: List[np.ndarray]) -> np.ndarray:
        s = np.zeros(len(lst) + 1, dtype=_INT)
        for i, a in enumerate(lst):
            s[i + 1] = s[i] + a.size
        return s

    def cat(lst: List[np.ndarray], dtype) -> np.ndarray:
        return np.concatenate([a.reshape(-1) for a in lst]).astype(dtype) if lst else np.zeros(0, dtype)

    gensegm_to_body = np.asarray(NBS.gensegm_to_body, dtype=_INT)
    gensegm_to_iint = np.asarray(NBS.gensegm_to_iint, dtype=_INT)
    inter_rot = np.zeros((nsegm, G, G))
    inter_isid = np.zeros(nsegm, dtype=_I32)
    inter_tr = np.zeros(nsegm, dtype=_INT)
    for isegm in range(nsegm):  # mirrors GatherInterSym, _NBodySyst.pyx:1282-1291
        Sym = NBS.intersegm_to_all[int(gensegm_to_body[isegm])][int(gensegm_to_iint[isegm])]
        inter_rot[isegm] = np.asarray(Sym.SpaceRot)
        inter_isid[isegm] = bool(Sym.IsIdentityRot())
        inter_tr[isegm] = int(Sym.TimeRev)
    assert (inter_tr == np.asarray(NBS.InterTimeRev)).all()

    loopgen = np.asarray(NBS.loopgen, dtype=_INT)
    ngensegm_loop = np.asarray(NBS.ngensegm_loop, dtype=_INT)
    alg_rot = np.zeros((nloop, G, G))
    for il in range(nloop):  # mirrors Compute_n_sub_fft, _NBodySyst.pyx:1391-1401
        Sym = NBS.gensegm_to_all[int(loopgen[il])][int(ngensegm_loop[il]) % nint_min]
        alg_rot[il] = np.asarray(Sym.SpaceRot)
        assert int(Sym.TimeRev) == int(NBS.ALG_TimeRev[il])

    shortcut = np.array([{"general_sym": GENERAL_SYM, "rfft": RFFT}[s] for s in NBS.ParamBasisShortcutPos],
                        dtype=_I32)
    law, law_params = _law_from_reference(NBS)

    return PlanTables(
        geodim=G, nbody=int(NBS.nbody), nloop=nloop, nsegm=nsegm, nbin=int(NBS.nbin_segm_unique),
        nint_min=nint_min, greater_nstore=bool(NBS.GreaterNStore),
        inter_law=law, inter_law_params=law_params,
        loopnb=np.asarray(NBS.loopnb), loopmass=np.asarray(NBS.loopmass),
        ncoeff_min=np.asarray(NBS.ncoeff_min_loop), n_sub_fft=np.asarray(NBS.n_sub_fft),
        shortcut=shortcut,
        nnz=np.array([b.shape[1] for b in basis_list], dtype=_INT),
        nppl=np.array([b.shape[2] for b in basis_list], dtype=_INT),
        nnz_k=cat(nnz_k_list, _INT), nnz_k_shifts=shifts(nnz_k_list),
        co_in=cat(co_in_list, _I32), co_in_shifts=shifts(co_in_list),
        params_basis=cat(basis_list, _C128), params_basis_shifts=shifts(basis_list),
        alg_iint=np.asarray(NBS.ALG_Iint), alg_timerev=np.asarray(NBS.ALG_TimeRev), alg_spacerot=alg_rot,
        bodyloop=np.asarray(NBS.bodyloop),
        bodysegm=np.asarray(NBS.bodysegm, dtype=_INT),
        body_spacerot=np.array([[np.asarray(NBS.intersegm_to_all[ib][ii].SpaceRot) for ii in range(nint_min)]
                                for ib in range(int(NBS.nbody))]),
        body_timerev=np.array([[int(NBS.intersegm_to_all[ib][ii].TimeRev) for ii in range(nint_min)]
                               for ib in range(int(NBS.nbody))], dtype=_INT),
        gensegm_to_body=gensegm_to_body, gensegm_to_iintrel=np.asarray(NBS.gensegm_to_iintrel),
        inter_timerev=inter_tr, inter_spacerot=inter_rot, inter_spacerot_isid=inter_isid,
        bin_source=np.asarray(NBS.BinSourceSegm), bin_target=np.asarray(NBS.BinTargetSegm),
        bin_spacerot=np.asarray(NBS.BinSpaceRot), bin_spacerot_isid=np.asarray(NBS.BinSpaceRotIsId),
        bin_prodchargesum=np.asarray(NBS.BinProdChargeSum),
    )
