"""choreographies2_b200 -- B200-native (sm_100a, fp64) action hot path of choreo's ``NBodySyst``.

Public surface: ``NBodySystB200`` (host mirror of the reference API for this path), ``PlanTables`` /
``extract_tables`` (symmetry tables read from a live reference object) and ``build`` (nvcc in-tree build).
"""
from .tables import PlanTables, extract_tables, GENERAL_SYM, RFFT  # noqa: F401
from .nbodysyst import NBodySystB200  # noqa: F401

__all__ = ["NBodySystB200", "PlanTables", "extract_tables", "GENERAL_SYM", "RFFT"]
