"""TEST INFRASTRUCTURE ONLY: compile the CTA programs of choreographies2_b200/csrc for the host.

The build container has no GPU.  ``g++ -DCHB_HOSTSIM`` turns every kernel launch of chb_api.cu into a loop over
blocks that runs the same CTA programs sequentially, so the index arithmetic, the plan logic and the Python host
layer can be checked against the oracle on the CPU.  The product never loads this library
(choreographies2_b200/_lib.py only loads libchoreo_b200.so); tests inject it explicitly.
"""
import ctypes
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
SRC = os.path.join(ROOT, "choreographies2_b200", "csrc", "chb_api.cu")
OUT = os.path.join(HERE, "libchb_hostsim.so")


def build(force=False):
    deps = [SRC, os.path.join(os.path.dirname(SRC), "chb_core.cuh"), os.path.join(os.path.dirname(SRC), "chb_fused.cuh"),
            os.path.join(os.path.dirname(SRC), "chb_fused4.cuh"),
            os.path.join(os.path.dirname(SRC), "chb_fused5.cuh"),
            os.path.join(os.path.dirname(SRC), "chb_big.cuh"),
            os.path.join(os.path.dirname(SRC), "chb_post.cuh"),
            os.path.join(os.path.dirname(SRC), "chb_newton.cuh"),
            os.path.join(os.path.dirname(SRC), "chb_krylov.cuh"),
            os.path.join(os.path.dirname(SRC), "chb_ode.cuh"),
            os.path.join(ROOT, "include", "choreo_b200.h")]
    if force or not os.path.isfile(OUT) or any(os.path.getmtime(d) > os.path.getmtime(OUT) for d in deps):
        subprocess.run(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-DCHB_HOSTSIM", "-x", "c++", SRC, "-o", OUT],
                       check=True)
    return OUT


def load():
    from choreographies2_b200 import _lib
    return _lib.bind(ctypes.CDLL(build()))
