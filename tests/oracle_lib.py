"""ctypes loader for oracle/libp3oracle.so — TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
The product package (scream_b200) never does.
"""
import ctypes
import os
import struct
import subprocess

import numpy as np

from scream_b200.fields import FIELDS, NUM_DIAG, RUNTIME_DEFAULTS

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
LIB_PATH = os.path.join(ORACLE_DIR, "libp3oracle.so")
TABLE_BLOB = os.path.join(ROOT, "scream_b200", "data", "p3_lookup_table_1.v4.1.1.f64")
REF_TABLE_TEXT = ("/root/reference/components/eam/src/physics/crm/pam/external/physics/micro/p3/tables/"
                  "p3_lookup_table_1.dat-v4.1.1")

_lib = None
_tables = None


def build(force=False):
    srcs = [os.path.join(ORACLE_DIR, f) for f in os.listdir(ORACLE_DIR) if f.endswith((".cpp", ".hpp"))]
    stale = (not os.path.exists(LIB_PATH)) or any(os.path.getmtime(s) > os.path.getmtime(LIB_PATH) for s in srcs)
    if force or stale:
        subprocess.check_call(["make", "-C", ORACLE_DIR, "-s"])


def lib():
    global _lib
    if _lib is None:
        build()
        L = ctypes.CDLL(LIB_PATH)
        L.p3o_read_ice_table_text.restype = ctypes.c_int
        L.p3o_read_ice_table_text.argtypes = [ctypes.c_char_p, ctypes.c_char_p, ctypes.c_void_p, ctypes.c_void_p]
        L.p3o_gen_rain_tables.restype = None
        L.p3o_gen_rain_tables.argtypes = [ctypes.c_void_p] * 4
        L.p3o_tables_create.restype = ctypes.c_void_p
        L.p3o_tables_create.argtypes = [ctypes.c_void_p] * 6
        L.p3o_tables_destroy.argtypes = [ctypes.c_void_p]
        L.p3o_main.restype = ctypes.c_double
        L.p3o_main.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p,
                               ctypes.c_int, ctypes.c_int, ctypes.c_double, ctypes.c_int, ctypes.c_int,
                               ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int]
        L.p3o_max_threads.restype = ctypes.c_int
        L.p3o_set_math_scale.restype = None
        L.p3o_set_math_scale.argtypes = [ctypes.c_void_p]
        L.p3o_set_math_mode.restype = None
        L.p3o_set_math_mode.argtypes = [ctypes.c_int]
        L.p3o_get_math_mode.restype = ctypes.c_int
        L.p3o_math_eval.restype = None
        L.p3o_math_eval.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
        _lib = L
    return _lib


def read_blob(path=TABLE_BLOB):
    raw = open(path, "rb").read()
    assert raw[:4] == b"P3TB"
    (n,) = struct.unpack("<I", raw[4:8])
    version = raw[8:8 + n].decode()
    body = np.frombuffer(raw, dtype="<f8", offset=8 + n)
    ice = body[:12000].reshape(5, 4, 50, 12).copy()
    coll = body[12000:72000].reshape(5, 4, 50, 30, 2).copy()
    return version, ice, coll


def read_text_table(path, version=b"4.1.1"):
    ice = np.zeros(12000)
    coll = np.zeros(60000)
    rc = lib().p3o_read_ice_table_text(path.encode(), version, ice.ctypes.data, coll.ctypes.data)
    if rc != 0:
        raise RuntimeError(f"p3o_read_ice_table_text rc={rc}")
    return ice.reshape(5, 4, 50, 12), coll.reshape(5, 4, 50, 30, 2)


def gen_rain_tables():
    vn, vm, revap, mu_r = np.zeros(3000), np.zeros(3000), np.zeros(3000), np.zeros(150)
    lib().p3o_gen_rain_tables(vn.ctypes.data, vm.ctypes.data, revap.ctypes.data, mu_r.ctypes.data)
    return vn.reshape(300, 10), vm.reshape(300, 10), revap.reshape(300, 10), mu_r


def tables():
    global _tables
    if _tables is None:
        _, ice, coll = read_blob()
        vn, vm, revap, mu_r = gen_rain_tables()
        h = lib().p3o_tables_create(ice.ctypes.data, coll.ctypes.data, vn.ctypes.data, vm.ctypes.data,
                                    revap.ctypes.data, mu_r.ctypes.data)
        _tables = (h, dict(ice=ice, coll=coll, vn=vn, vm=vm, revap=revap, mu_r=mu_r))
    return _tables


def runtime_array(overrides=None):
    vals = dict(RUNTIME_DEFAULTS)
    if overrides:
        vals.update(overrides)
    return np.array([float(vals[k]) for k, _ in RUNTIME_DEFAULTS], dtype=np.float64)


def p3_main(fields, dt, predict_nc=True, prescribed_ccn=True, runtime=None, nthreads=0, want_diag=True):
    """Run the oracle in place on a dict of numpy arrays (see scream_b200.fields.FIELDS).

    Returns (elapsed_seconds, diag[ncol, NUM_DIAG] or None)."""
    qc = fields["qc"]
    ncol, nlev = qc.shape
    dtype = qc.dtype
    real_bytes = dtype.itemsize
    ptrs = (ctypes.c_void_p * len(FIELDS))()
    strides = (ctypes.c_int64 * len(FIELDS))()
    for i, n in enumerate(FIELDS):
        a = fields[n]
        assert a.dtype == dtype and a.flags["C_CONTIGUOUS"] or a.ndim == 2, n
        assert a.dtype == dtype, n
        ptrs[i] = a.ctypes.data
        strides[i] = (a.strides[0] // real_bytes) if a.ndim == 2 else 1
        if a.ndim == 2:
            assert a.strides[1] == real_bytes, n
    diag = np.full((ncol, NUM_DIAG), -7, dtype=np.int32) if want_diag else None
    rt = runtime_array(runtime)
    h, _ = tables()
    t = lib().p3o_main(h, real_bytes, ptrs, strides, ncol, nlev, float(dt), int(predict_nc), int(prescribed_ccn),
                       rt.ctypes.data, diag.ctypes.data if want_diag else None, nthreads)
    if t < 0:
        raise RuntimeError(f"p3o_main failed rc={t}")
    return t, diag


MATH_FUNCS = ("pow", "exp", "log", "log10", "cbrt", "tgamma", "tanh", "expm1")


def set_math_scale(scale=None):
    """Conditioning probe (oracle/p3_oracle_math.hpp): multiply the result of every transcendental call by
    scale[fn] (8 values in MATH_FUNCS order, or one value for all); None restores exact libm."""
    if scale is None:
        lib().p3o_set_math_scale(None)
        return
    arr = np.full(len(MATH_FUNCS), scale, dtype=np.float64) if np.isscalar(scale) else np.asarray(scale, dtype=np.float64)
    assert arr.shape == (len(MATH_FUNCS),)
    lib().p3o_set_math_scale(arr.ctypes.data)


def set_math_mode(mode):
    """'libm': the host's std:: math (what the reference's CPU build calls); 'shared': the deterministic
    transcendentals of scream_b200/csrc/p3_math.hpp that the kernels are built with (-DP3B200_SHARED_MATH)."""
    assert mode in ("libm", "shared")
    lib().p3o_set_math_mode(1 if mode == "shared" else 0)


def math_mode():
    return "shared" if lib().p3o_get_math_mode() else "libm"


def math_eval(fn, x, y=None, real_bytes=8):
    """Evaluate transcendental `fn` (a name in MATH_FUNCS) of the shared header on arrays."""
    x = np.ascontiguousarray(x, dtype=np.float64)
    out = np.empty_like(x)
    yy = np.ascontiguousarray(y, dtype=np.float64) if y is not None else None
    lib().p3o_math_eval(MATH_FUNCS.index(fn), real_bytes, x.size, x.ctypes.data,
                        yy.ctypes.data if yy is not None else None, out.ctypes.data)
    return out


def max_threads():
    return lib().p3o_max_threads()


def _prepost_ptrs(fields, names, optional=()):
    ptrs = (ctypes.c_void_p * len(names))()
    for i, n in enumerate(names):
        a = fields.get(n)
        if a is None:
            assert n in optional, n
            ptrs[i] = None
        else:
            assert a.flags["C_CONTIGUOUS"], n
            ptrs[i] = a.ctypes.data
    return ptrs


def preamble(fields):
    """p3_preamble restatement (oracle/p3_oracle_prepost.cpp) in place on numpy arrays keyed by PREAMBLE_FIELDS."""
    from scream_b200.fields import PREAMBLE_FIELDS
    L = lib()
    ncol, nlev = fields["qv"].shape
    rc = L.p3o_preamble(fields["qv"].dtype.itemsize, _prepost_ptrs(fields, PREAMBLE_FIELDS), ncol, nlev)
    assert rc == 0


def postamble(fields, dt):
    """p3_postamble restatement in place; the four boundary-flux arrays are optional."""
    from scream_b200.fields import POSTAMBLE_FIELDS_1D, POSTAMBLE_FIELDS_2D
    L = lib()
    ncol, nlev = fields["qv"].shape
    names = POSTAMBLE_FIELDS_2D + POSTAMBLE_FIELDS_1D
    rc = L.p3o_postamble(fields["qv"].dtype.itemsize, _prepost_ptrs(fields, names, optional=POSTAMBLE_FIELDS_1D[4:]),
                         ncol, nlev, ctypes.c_double(dt))
    assert rc == 0
