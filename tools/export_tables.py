#!/usr/bin/env python
"""Export the symmetry tables of the BASELINE.json systems into choreographies2_b200/systems/*.npz.

Symmetry setup is out of scope for the B200 path: the tables are produced once, here, by the
unmodified reference (oracle/_ref, built by oracle/build_ref.sh) and committed, so that bench.py, smoke() and the GPU
tests need neither /root/reference nor the reference package at run time.

    python tools/export_tables.py
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import ref_loader  # noqa: E402
from choreographies2_b200.tables import extract_tables  # noqa: E402

OUT = os.path.join(ROOT, "choreographies2_b200", "systems")
NAMES = ["3C", "3B", "5Dq", "20B", "16C", "64C", "3D", "3C5k"]

if __name__ == "__main__":
    systems = ref_loader.load_systems()
    for name in NAMES:
        NBS = ref_loader.build_system(systems[name])
        T = extract_tables(NBS)
        T.save(os.path.join(OUT, f"{name}.npz"))
        print(name, "nloop", T.nloop, "nsegm", T.nsegm, "nbin", T.nbin, "nint_min", T.nint_min)
