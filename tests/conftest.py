import glob
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")

# Parity tolerances (BASELINE.json north_star): action & gradient <= 1e-11, Hessian-vector <= 1e-10,
# relative to the largest entry of the reference result.
TOL_ACTION = 1e-11
TOL_GRAD = 1e-11
TOL_HESS = 1e-10
TOL_TRANSFORM = 1e-11


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def golden_names():
    return sorted(os.path.basename(p)[len("golden_"):-len(".npz")]
                  for p in glob.glob(os.path.join(GOLDEN_DIR, "golden_*.npz")))


class Golden:
    """One golden_<name>.npz: variants -> tables + per-nint vectors."""

    def __init__(self, name):
        self.name = name
        self.z = np.load(os.path.join(GOLDEN_DIR, f"golden_{name}.npz"))
        self.variants = {}
        for k in self.z.files:
            v, lvl, _ = k.split("/", 2)
            self.variants.setdefault(v, set())
            if lvl != "T":
                self.variants[v].add(int(lvl))

    def tables_dict(self, variant):
        pre = f"{variant}/T/"
        return {k[len(pre):]: self.z[k] for k in self.z.files if k.startswith(pre)}

    def vec(self, variant, nint, key):
        return self.z[f"{variant}/{nint}/{key}"]

    def cases(self):
        for v in sorted(self.variants):
            for nint in sorted(self.variants[v]):
                yield v, nint


@pytest.fixture(scope="session")
def golden_cache():
    cache = {}

    def get(name):
        if name not in cache:
            cache[name] = Golden(name)
        return cache[name]

    return get


def relerr(a, b):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    if a.size == 0:
        return 0.0
    den = max(float(np.max(np.abs(b))), 1e-300)
    return float(np.max(np.abs(a - b))) / den


@pytest.fixture(params=["hostsim", pytest.param("cuda", marks=pytest.mark.gpu)])
def backend(request):
    """Library handle a reference-comparison test runs on: the g++ emulation of the CTA programs (CPU suite) or the CUDA
    library (``-m gpu``; ``None`` selects the product loader).  The reference (oracle/_ref) is the checker in both."""
    if request.param == "hostsim":
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        from hostsim import build_hostsim
        return build_hostsim.load()
    from choreographies2_b200 import _lib
    assert _lib.load().chb_device_count() >= 1, "no sm_100 device visible: GPU tests need a B200"
    return None


def make_nbs(tables, nint, backend):
    from choreographies2_b200 import NBodySystB200
    return NBodySystB200(tables, nint) if backend is None else NBodySystB200(tables, nint, _lib_handle=backend)


def need_reference():
    from oracle import ref_loader
    if not ref_loader.reference_available():
        pytest.skip("oracle/_ref not built")
    ref_loader.import_reference()
    return ref_loader
