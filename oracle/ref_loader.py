"""TEST INFRASTRUCTURE ONLY -- access to the UNMODIFIED reference built into ``oracle/_ref``.

``import_reference()`` applies the two shims SURVEY.md section 8(c) lists (matplotlib stub lives
in ``oracle/_ref/matplotlib``; ``scipy.optimize.nonlin`` alias for scipy >= 1.14) and imports
``choreo`` from ``oracle/_ref``.  ``build_system(desc)`` instantiates ``choreo.NBodySyst`` from a
plain-JSON system description (the entries of ``tests/golden/systems.json``, which were exported
from the reference's own ``tests/NewSym_data/*/choreo_config.json`` by ``tests/golden/make_golden.py``),
so nothing here needs ``/root/reference`` at run time.
"""
from __future__ import annotations

import json
import os
import sys
from typing import Any, Dict, Optional

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(_HERE, "_ref")
SYSTEMS_JSON = os.path.join(os.path.dirname(_HERE), "tests", "golden", "systems.json")

_choreo = None


def reference_available() -> bool:
    return os.path.isfile(os.path.join(REF_DIR, ".built"))


def import_reference():
    """Import the reference ``choreo`` package from ``oracle/_ref`` (raises ImportError if not built)."""
    global _choreo
    if _choreo is not None:
        return _choreo
    if not reference_available():
        raise ImportError("oracle/_ref is not built: run oracle/build_ref.sh (needs /root/reference)")
    import scipy.optimize  # noqa: F401
    try:
        import scipy.optimize.nonlin  # noqa: F401
    except Exception:
        pass
    import scipy.optimize._nonlin as _nl
    # choreo/scipy_plus/nonlin.py:78-89 needs names that scipy>=1.14 only keeps in the private module
    scipy.optimize.nonlin = _nl
    sys.modules["scipy.optimize.nonlin"] = _nl
    if REF_DIR not in sys.path:
        sys.path.insert(0, REF_DIR)
    import choreo  # type: ignore

    assert os.path.abspath(choreo.__file__).startswith(REF_DIR), choreo.__file__
    _choreo = choreo
    return choreo


def load_systems() -> Dict[str, Any]:
    with open(SYSTEMS_JSON) as f:
        return json.load(f)


def sym_list_from_desc(desc: Dict[str, Any]):
    choreo = import_reference()
    out = []
    for s in desc["syms"]:
        out.append(choreo.ActionSym(
            np.array(s["BodyPerm"], dtype=np.intp),
            np.array(s["SpaceRot"], dtype=np.float64),
            int(s["TimeRev"]), int(s["TimeShiftNum"]), int(s["TimeShiftDen"]),
        ))
    return out


def build_system(desc: Dict[str, Any], inter_pow: Optional[float] = None, ForceGeneralSym: bool = False,
                 ForceGreaterNStore: bool = False):
    """``choreo.NBodySyst`` for a system description, same law selection as the reference tests
    (tests/test_config/test_config_NBodySyst.py:47-77)."""
    choreo = import_reference()
    ipow = desc["inter_pow"] if inter_pow is None else inter_pow
    ipm = desc["inter_pm"]
    if ipow == -1.0 and ipm == 1:
        law, law_par = "gravity_pot", None
    else:
        law, law_par = "power_law_pot", {"n": ipow, "alpha": ipm}
    return choreo.NBodySyst(
        int(desc["geodim"]), int(desc["nbody"]),
        np.array(desc["mass"], dtype=np.float64), np.array(desc["charge"], dtype=np.float64),
        sym_list_from_desc(desc), inter_law_str=law, inter_law_param_dict=law_par,
        ForceGeneralSym=ForceGeneralSym, ForceGreaterNStore=ForceGreaterNStore,
    )


def build_named(name: str, **kw):
    return build_system(load_systems()[name], **kw)
