"""CPU tests of the C-ABI boundary: the CUDA library builds, loads and exports every symbol that
include/choreo_b200.h declares; without a GPU the product path fails loudly (no CPU fallback)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from conftest import ROOT
from choreographies2_b200 import _lib, build as chb_build
from choreographies2_b200.tables import PlanTables


@pytest.fixture(scope="module")
def cuda_lib():
    chb_build.build()
    return _lib.load()


def _declared_symbols():
    hdr = open(os.path.join(ROOT, "include", "choreo_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    return sorted(set(re.findall(r"\b(chb_[a-zA-Z0-9_]+)\s*\(", hdr)))


def test_header_and_binding_agree(cuda_lib):
    declared = _declared_symbols()
    assert len(declared) >= 30
    assert sorted(_lib.SYMBOLS) == declared
    for name in declared:
        assert hasattr(cuda_lib, name), name


def test_library_is_sm100a_only():
    import subprocess
    out = subprocess.run(["cuobjdump", "--list-elf", _lib.LIB_PATH], capture_output=True, text=True).stdout
    assert "sm_100a" in out
    assert not re.search(r"sm_(?!100a)\d+", out)


def test_abi_version_and_errors(cuda_lib):
    assert cuda_lib.chb_abi_version() == 2
    plan = C.c_void_p()
    assert cuda_lib.chb_plan_create(None, 0, C.byref(plan)) == 1  # CHB_ERR_INVALID_ARGUMENT
    assert b"NULL" in cuda_lib.chb_last_error_string()


def test_no_cpu_fallback_without_gpu(cuda_lib):
    """On a box without a CUDA device plan creation must fail with CHB_ERR_NO_DEVICE."""
    if cuda_lib.chb_device_count() > 0:
        pytest.skip("a B200 is visible")
    T = PlanTables.load(os.path.join(ROOT, "choreographies2_b200", "systems", "3C.npz"))
    from choreographies2_b200 import NBodySystB200
    with pytest.raises(_lib.ChoreoB200Error, match="CHB_ERR_NO_DEVICE"):
        NBodySystB200(T, 24)


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "choreographies2_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in src and "from oracle" not in src, f
                # the CTA programs under csrc/ are also compiled by tests/hostsim (test-only emulation) and say so in
                # comments; nothing on the Python side of the product may know about it
                assert "hostsim" not in src or f.endswith((".cu", ".cuh")), f
