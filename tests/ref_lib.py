"""ctypes loader for oracle/_ref/libp3ref_*.so — the reference's OWN p3_main compiled against the Kokkos stand-in
(oracle/Makefile.ref, oracle/ref_shim/).  TEST INFRASTRUCTURE ONLY: used to pin the oracle restatement
(tests/test_cpu_oracle_vs_ref.py), to generate the committed fixtures (tools/make_ref_golden.py) and by bench.py's
CPU-baseline legs.  The product package never imports this.

The libraries can only be BUILT where /root/reference exists (this container); the prebuilt .so files travel to the
GPU box with the repository snapshot.
"""
import ctypes
import os
import subprocess

import numpy as np

from scream_b200.fields import FIELDS
from tests import oracle_lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
REF_DIR = os.path.join(ORACLE_DIR, "_ref")
REFERENCE_TREE = "/root/reference"

_libs = {}
_tables = {}


HOOK = "hook"   # pseudo pack size: the pack-size-1 library with the switchable math symbols (shared-math mode)


def lib_path(real_bytes=8, pack=1):
    suffix = "pk1_hook" if pack == HOOK else f"pk{pack}"
    return os.path.join(REF_DIR, f"libp3ref_{'dp' if real_bytes == 8 else 'sp'}_{suffix}.so")


def can_build():
    return os.path.isdir(os.path.join(REFERENCE_TREE, "components", "eamxx", "src", "physics", "p3"))


def build():
    """Compile oracle/_ref from the reference sources in place (only possible where /root/reference exists)."""
    if can_build():
        subprocess.check_call(["make", "-C", ORACLE_DIR, "-f", "Makefile.ref", "-s", "-j4"])


def available(real_bytes=8, pack=1):
    return os.path.exists(lib_path(real_bytes, pack))


def lib(real_bytes=8, pack=1):
    key = (real_bytes, pack)
    if key not in _libs:
        if not available(real_bytes, pack):
            build()
        L = ctypes.CDLL(lib_path(real_bytes, pack))
        L.p3ref_real_bytes.restype = ctypes.c_int
        L.p3ref_pack_size.restype = ctypes.c_int
        L.p3ref_max_threads.restype = ctypes.c_int
        L.p3ref_tables_create.restype = ctypes.c_void_p
        L.p3ref_tables_create.argtypes = [ctypes.c_void_p] * 6
        L.p3ref_tables_load_ice_with_reference_loader.restype = ctypes.c_int
        L.p3ref_tables_load_ice_with_reference_loader.argtypes = [ctypes.c_void_p] * 3
        L.p3ref_tables_destroy.argtypes = [ctypes.c_void_p]
        L.p3ref_main.restype = ctypes.c_double
        L.p3ref_main.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p,
                                 ctypes.c_int, ctypes.c_int, ctypes.c_double, ctypes.c_int, ctypes.c_int,
                                 ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int]
        assert L.p3ref_real_bytes() == real_bytes and L.p3ref_pack_size() == (1 if pack == HOOK else pack)
        if pack == HOOK:
            L.p3ref_set_math_mode.restype = None
            L.p3ref_set_math_mode.argtypes = [ctypes.c_int]
            L.p3ref_get_math_mode.restype = ctypes.c_int
        _libs[key] = L
    return _libs[key]


def tables(real_bytes=8, pack=1):
    """Lookup tables of the reference build: ice/collection values from the repository's blob (compared with the
    reference's own text loader in tests/test_cpu_oracle_vs_ref.py), rain tables from the restated generator."""
    key = (real_bytes, pack)
    if key not in _tables:
        _, t = oracle_lib.tables()
        keep = {k: np.ascontiguousarray(v, dtype=np.float64) for k, v in t.items()}
        h = lib(real_bytes, pack).p3ref_tables_create(keep["ice"].ctypes.data, keep["coll"].ctypes.data,
                                                      keep["vn"].ctypes.data, keep["vm"].ctypes.data,
                                                      keep["revap"].ctypes.data, keep["mu_r"].ctypes.data)
        _tables[key] = (h, keep)
    return _tables[key]


def load_ice_tables_with_reference_loader(real_bytes=8, pack=1):
    """Run the reference's init_kokkos_ice_lookup_tables on the reference's own table file; returns (ice, log10 coll)."""
    h, _ = tables(real_bytes, pack)
    ice = np.zeros(12000)
    coll = np.zeros(60000)
    rc = lib(real_bytes, pack).p3ref_tables_load_ice_with_reference_loader(h, ice.ctypes.data, coll.ctypes.data)
    if rc != 0:
        raise RuntimeError("reference table loader failed")
    return ice.reshape(5, 4, 50, 12), coll.reshape(5, 4, 50, 30, 2)


def p3_main(fields, dt, predict_nc=True, prescribed_ccn=True, runtime=None, nthreads=0, pack=1, math="libm"):
    """Run the reference's p3_main in place on a dict of numpy arrays (scream_b200.fields.FIELDS).
    math="shared" runs the hooked library with the transcendentals of scream_b200/csrc/p3_math.hpp (the arithmetic of
    the GPU parity build and of oracle_lib.set_math_mode("shared")); math="libm" is the reference as it is.
    Returns the elapsed seconds the reference itself measured (p3_main_impl.hpp:105,326-330)."""
    assert math in ("libm", "shared")
    if math == "shared":
        pack = HOOK
    qc = fields["qc"]
    ncol, nlev = qc.shape
    dtype = qc.dtype
    real_bytes = dtype.itemsize
    ptrs = (ctypes.c_void_p * len(FIELDS))()
    strides = (ctypes.c_int64 * len(FIELDS))()
    for i, n in enumerate(FIELDS):
        a = fields[n]
        assert a.dtype == dtype, n
        ptrs[i] = a.ctypes.data
        strides[i] = (a.strides[0] // real_bytes) if a.ndim == 2 else 1
        if a.ndim == 2:
            assert a.strides[1] == real_bytes, n
    rt = oracle_lib.runtime_array(runtime)
    h, _ = tables(real_bytes, pack)      # tables are always built in libm mode (log10 of the collection table)
    L = lib(real_bytes, pack)
    if pack == HOOK:
        L.p3ref_set_math_mode(1 if math == "shared" else 0)
    try:
        t = L.p3ref_main(h, real_bytes, ptrs, strides, ncol, nlev, float(dt), int(predict_nc),
                         int(prescribed_ccn), rt.ctypes.data, None, nthreads)
    finally:
        if pack == HOOK:
            L.p3ref_set_math_mode(0)
    if t < 0:
        raise RuntimeError(f"p3ref_main failed rc={t}")
    return t


def max_threads(real_bytes=8, pack=1):
    return lib(real_bytes, pack).p3ref_max_threads()
