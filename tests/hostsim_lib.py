"""ctypes loader for the 1-lane host build of the column driver (tests/hostsim) — TEST INFRASTRUCTURE ONLY.

The hostsim library compiles scream_b200/csrc/p3_column.cuh + p3_physics.cuh (the exact source the
sm_100a kernel instantiates) with g++ and a warp of size 1, so that the driver's control flow, level
indexing, the register-window sedimentation and the host table code can be compared with the oracle
on a machine without a GPU.  Nothing in the product package imports it.
"""
import ctypes
import os
import subprocess

import numpy as np

from scream_b200.fields import FIELDS, NUM_DIAG
from scream_b200 import lib as p3

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = [os.path.join(ROOT, "tests", "hostsim", "p3_hostsim.cpp"),
       os.path.join(ROOT, "scream_b200", "csrc", "p3_host.cpp")]
DEPS = SRC + [os.path.join(ROOT, "scream_b200", "csrc", f) for f in
              ("p3_column.cuh", "p3_physics.cuh", "p3_host.hpp", "p3_math.hpp")] + [os.path.join(ROOT, "include", "p3_b200.h")]
# two flavours: "exact" (-DP3B200_EXACT_MATH: the reference's own pow calls, comparable with the oracle to
# the last bit -> proves the driver logic) and "product" (the arithmetic the GPU build uses, which replaces a
# few pow calls by 1-ulp-equivalent cheaper forms -> compared on the well-conditioned columns)
# "parity" (-DP3B200_PARITY = exact order + the shared deterministic transcendentals of p3_math.hpp): has to match
# the oracle in shared-math mode BIT FOR BIT on every column; it is the CPU twin of libp3b200_parity.so
LIB_PATHS = {"exact": os.environ.get("P3B200_HOSTSIM_EXACT_LIB", os.path.join(ROOT, "build", "libp3hostsim_exact.so")),   # override: tools/hostsim_sanitize.sh
             "product": os.path.join(ROOT, "build", "libp3hostsim.so"),
             "parity": os.path.join(ROOT, "build", "libp3hostsim_parity.so")}

_libs = {}
_tables = None


def build(flavour="exact"):
    path = LIB_PATHS[flavour]
    stale = (not os.path.exists(path)) or any(os.path.getmtime(s) > os.path.getmtime(path) for s in DEPS)
    if stale and not (flavour == "exact" and "P3B200_HOSTSIM_EXACT_LIB" in os.environ):
        os.makedirs(os.path.dirname(path), exist_ok=True)
        extra = {"exact": ["-DP3B200_EXACT_MATH"], "parity": ["-DP3B200_PARITY"], "product": []}[flavour]
        subprocess.check_call(["g++", "-O2", "-march=x86-64-v3", "-std=c++17", "-fPIC", "-ffp-contract=off", "-shared",
                               "-I/usr/local/cuda/include"] + extra + ["-o", path] + SRC)


def lib(flavour="exact"):
    if flavour not in _libs:
        build(flavour)
        L = ctypes.CDLL(LIB_PATHS[flavour])
        L.p3sim_tables.argtypes = [ctypes.c_char_p, ctypes.POINTER(p3.P3B200Tables)]
        L.p3sim_main.argtypes = [ctypes.POINTER(p3.P3B200Tables), ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p,
                                 ctypes.c_int, ctypes.c_int, ctypes.c_double, ctypes.c_int, ctypes.c_int,
                                 ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]
        _libs[flavour] = L
    return _libs[flavour]


def tables():
    global _tables
    if _tables is None:
        t = p3.P3B200Tables()
        rc = lib().p3sim_tables(p3.TABLE_BLOB.encode(), ctypes.byref(t))
        assert rc == 0, rc
        _tables = t
    return _tables


def p3_main(fields, dt, predict_nc=True, prescribed_ccn=True, zero_flux_top=True, flavour="exact"):
    """Run the 1-lane host build in place on a dict of numpy arrays; returns diag[ncol, NUM_DIAG]."""
    qc = fields["qc"]
    ncol, nlev = qc.shape
    rb = qc.dtype.itemsize
    ptrs = (ctypes.c_void_p * len(FIELDS))()
    strides = (ctypes.c_int64 * len(FIELDS))()
    for i, n in enumerate(FIELDS):
        a = fields[n]
        assert a.dtype == qc.dtype, n
        ptrs[i] = a.ctypes.data
        strides[i] = (a.strides[0] // rb) if a.ndim == 2 else 1
    diag = np.full((ncol, NUM_DIAG), -7, dtype=np.int32)
    rc = lib(flavour).p3sim_main(ctypes.byref(tables()), rb, ptrs, strides, ncol, nlev, float(dt), int(predict_nc),
                          int(prescribed_ccn), int(zero_flux_top), None, diag.ctypes.data)
    assert rc == 0, rc
    return diag
