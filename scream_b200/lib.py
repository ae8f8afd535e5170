"""ctypes binding of the product library libp3b200.so (C ABI: include/p3_b200.h).

This is plumbing only: torch supplies device memory and streams, numpy supplies host arrays; every
P3 computation happens in the sm_100a kernel behind p3b200_main / p3b200_main_host.  There is no CPU
fallback: importing works without a GPU (so the ABI can be inspected), creating a handle does not.
"""
import ctypes
import os
import subprocess

import numpy as np

from .fields import (CONSERVATION_FIELDS, CHECK_COLUMN_FIELDS, CHECK_FIELDS, FIELDS, INTERFACE_FIELDS, NUM_DIAG, POSTAMBLE_FIELDS_1D, POSTAMBLE_FIELDS_2D, PREAMBLE_FIELDS,
                     RUNTIME_DEFAULTS, SURFACE_1D)

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("P3B200_LIB", os.path.join(PKG_DIR, "libp3b200.so"))   # override: development only
# Build flavours of the same sources (scream_b200/csrc/Makefile): "product" is what ships; "parity" is compiled with
# -DP3B200_PARITY (the reference's operation order everywhere + the deterministic transcendentals of p3_math.hpp) and
# has to agree with the shared-math oracle bit for bit on every column (tests/test_gpu_parity_all_columns.py).
LIB_PATHS = {"product": LIB_PATH, "parity": os.path.join(PKG_DIR, "libp3b200_parity.so")}
TABLE_BLOB = os.path.join(PKG_DIR, "data", "p3_lookup_table_1.v4.1.1.f64")
NUM_FIELDS = len(FIELDS)
STATIC_BUDGET_MAX_COLS = 32768   # P3B200_STATIC_BUDGET_MAX_COLS (include/p3_b200.h)

# every symbol include/p3_b200.h declares
ABI_SYMBOLS = [
    "p3b200_abi_version", "p3b200_strerror", "p3b200_last_cuda_error", "p3b200_device_count",
    "p3b200_runtime_defaults", "p3b200_tables_read_text", "p3b200_tables_read_blob",
    "p3b200_tables_generate_rain", "p3b200_create", "p3b200_destroy", "p3b200_main", "p3b200_get_diag",
    "p3b200_error_summary", "p3b200_main_host", "p3b200_last_kernel_ms", "p3b200_last_phase_ms",
    "p3b200_last_sed_concurrent", "p3b200_last_phase_a_ms",
    "p3b200_main_preamble", "p3b200_column_conservation_begin", "p3b200_column_conservation_check", "p3b200_column_conservation_get",
    "p3b200_water_energy_totals", "p3b200_water_energy_totals_static", "p3b200_set_default_tables", "p3b200_preamble", "p3b200_postamble",
    "p3b200_property_checks", "p3b200_postamble_checked", "p3b200_repair_qv", "p3b200_main_postamble",
    "p3b200_p3_main_f",
]


class P3B200Tables(ctypes.Structure):
    _fields_ = [("mu_r_table_vals", ctypes.c_double * 150),
                ("vn_table_vals", ctypes.c_double * 3000),
                ("vm_table_vals", ctypes.c_double * 3000),
                ("revap_table_vals", ctypes.c_double * 3000),
                ("ice_table_vals", ctypes.c_double * 12000),
                ("collect_table_vals", ctypes.c_double * 60000),
                ("dnu_table_vals", ctypes.c_double * 16)]


class P3B200Runtime(ctypes.Structure):
    _fields_ = [(name, ctypes.c_double) for name, _ in RUNTIME_DEFAULTS[:-1]] + \
               [("do_ice_production", ctypes.c_int32), ("pad_", ctypes.c_int32)]


class P3B200Fields(ctypes.Structure):
    _fields_ = [("ptr", ctypes.c_void_p * NUM_FIELDS), ("col_stride", ctypes.c_int64 * NUM_FIELDS)]


class P3B200Infra(ctypes.Structure):
    _fields_ = [("dt", ctypes.c_double), ("it", ctypes.c_int32), ("ncol", ctypes.c_int32),
                ("nlev", ctypes.c_int32), ("predictNc", ctypes.c_int32), ("prescribedCCN", ctypes.c_int32),
                ("zero_flux_top", ctypes.c_int32), ("zero_history", ctypes.c_int32)]


class P3Error(RuntimeError):
    pass


_libs = {}


def build(force=False):
    """Compile libp3b200.so in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    csrc = os.path.join(PKG_DIR, "csrc")
    if force:
        subprocess.check_call(["make", "-C", csrc, "-s", "clean"])
    subprocess.check_call(["make", "-C", csrc, "-s"])


def lib(flavour="product"):
    if flavour not in _libs:
        path = LIB_PATHS[flavour]
        if not os.path.exists(path):
            raise P3Error(f"{path} is missing: build it with `make -C scream_b200/csrc` "
                          "(there is no CPU fallback for p3_main)")
        L = ctypes.CDLL(path)
        L.p3b200_strerror.restype = ctypes.c_char_p
        L.p3b200_last_cuda_error.restype = ctypes.c_char_p
        L.p3b200_runtime_defaults.restype = None
        L.p3b200_create.argtypes = [ctypes.POINTER(P3B200Tables), ctypes.c_int, ctypes.c_int,
                                    ctypes.POINTER(ctypes.c_void_p)]
        L.p3b200_destroy.argtypes = [ctypes.c_void_p]
        L.p3b200_main.argtypes = [ctypes.c_void_p, ctypes.POINTER(P3B200Runtime), ctypes.POINTER(P3B200Fields),
                                  ctypes.POINTER(P3B200Infra), ctypes.c_void_p,
                                  ctypes.POINTER(ctypes.c_int64), ctypes.c_void_p]
        L.p3b200_main_host.argtypes = [ctypes.c_void_p, ctypes.POINTER(P3B200Runtime),
                                       ctypes.POINTER(P3B200Fields), ctypes.POINTER(P3B200Infra),
                                       ctypes.POINTER(ctypes.c_int64), ctypes.c_void_p]
        L.p3b200_get_diag.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int32]
        L.p3b200_error_summary.argtypes = [ctypes.c_void_p, ctypes.POINTER(ctypes.c_int32)]
        L.p3b200_last_kernel_ms.argtypes = [ctypes.c_void_p, ctypes.POINTER(ctypes.c_double),
                                            ctypes.POINTER(ctypes.c_int64)]
        L.p3b200_last_phase_ms.argtypes = [ctypes.c_void_p, ctypes.POINTER(ctypes.c_double)]
        L.p3b200_last_sed_concurrent.argtypes = [ctypes.c_void_p]
        L.p3b200_last_phase_a_ms.argtypes = [ctypes.c_void_p, ctypes.POINTER(ctypes.c_double)]
        L.p3b200_main_preamble.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                           ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
        L.p3b200_column_conservation_begin.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int32, ctypes.c_int32, ctypes.c_void_p]
        L.p3b200_column_conservation_check.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int32, ctypes.c_int32,
                                                       ctypes.c_double, ctypes.c_double, ctypes.c_double, ctypes.c_void_p,
                                                       ctypes.c_void_p]
        L.p3b200_column_conservation_get.argtypes = [ctypes.c_void_p, ctypes.c_int32] + [ctypes.c_void_p] * 4
        L.p3b200_water_energy_totals.argtypes = [ctypes.c_void_p, ctypes.POINTER(P3B200Fields), ctypes.POINTER(P3B200Infra),
                                                 ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
        L.p3b200_water_energy_totals_static.argtypes = [ctypes.c_void_p, ctypes.POINTER(P3B200Fields),
                                                        ctypes.POINTER(P3B200Infra), ctypes.c_void_p, ctypes.c_void_p]
        L.p3b200_set_default_tables.argtypes = [ctypes.POINTER(P3B200Tables), ctypes.c_int]
        L.p3b200_tables_read_text.argtypes = [ctypes.c_char_p, ctypes.c_char_p, ctypes.POINTER(P3B200Tables)]
        L.p3b200_tables_read_blob.argtypes = [ctypes.c_char_p, ctypes.c_char_p, ctypes.POINTER(P3B200Tables)]
        L.p3b200_tables_generate_rain.argtypes = [ctypes.POINTER(P3B200Tables)]
        dp = ctypes.POINTER(ctypes.c_double)
        L.p3b200_p3_main_f.argtypes = ([dp] * 6 + [ctypes.c_double] + [dp] * 10 + [ctypes.c_int] + [dp] * 2 +
                                       [ctypes.c_int] * 4 + [dp] * 4 + [ctypes.c_bool] * 2 + [dp] * 13)
        _libs[flavour] = L
    return _libs[flavour]


def check(rc, what="p3b200"):
    if rc != 0:
        L = lib()
        msg = L.p3b200_strerror(rc).decode()
        cuda = L.p3b200_last_cuda_error().decode()
        raise P3Error(f"{what} failed: {msg} (rc={rc}) {cuda}")


_tables_cache = None


def load_tables(blob=TABLE_BLOB, version=b"4.1.1"):
    """Host-side P3LookupTables: ice/collection from the shipped blob, rain tables generated
    (init_kokkos_ice_lookup_tables + init_kokkos_tables + p3_init_b)."""
    global _tables_cache
    if _tables_cache is None or blob != TABLE_BLOB:
        t = P3B200Tables()
        check(lib().p3b200_tables_read_blob(blob.encode(), version, ctypes.byref(t)), "tables_read_blob")
        check(lib().p3b200_tables_generate_rain(ctypes.byref(t)), "tables_generate_rain")
        if blob != TABLE_BLOB:
            return t
        _tables_cache = t
    return _tables_cache


class P3B200PreambleFields(ctypes.Structure):
    _fields_ = [(n, ctypes.c_void_p) for n in PREAMBLE_FIELDS] + [("col_stride", ctypes.c_int64)]


class P3B200PostambleFields(ctypes.Structure):
    # member order of include/p3_b200.h
    _fields_ = [(n, ctypes.c_void_p) for n in
                ("pmid", "pseudo_density", "pseudo_density_dry", "th_atm", "T_atm",
                 "qv", "qc", "nc", "qr", "nr", "qi", "qm", "ni", "bm",
                 "diag_eff_radius_qc", "diag_eff_radius_qi", "diag_eff_radius_qr", "T_prev", "qv_prev",
                 "precip_liq_surf_flux", "precip_ice_surf_flux", "precip_liq_surf_mass", "precip_ice_surf_mass",
                 "vapor_flux", "water_flux", "ice_flux", "heat_flux")] + [("col_stride", ctypes.c_int64), ("dt", ctypes.c_double)]


class P3B200CheckFields(ctypes.Structure):
    _fields_ = [(n, ctypes.c_void_p) for n in CHECK_FIELDS] + [("col_stride", ctypes.c_int64)]


class P3B200CheckResult(ctypes.Structure):
    _fields_ = [("lb", ctypes.c_double), ("ub", ctypes.c_double), ("min_val", ctypes.c_double), ("max_val", ctypes.c_double),
                ("min_loc", ctypes.c_int64), ("max_loc", ctypes.c_int64), ("result", ctypes.c_int32),
                ("pass_lower", ctypes.c_int32), ("pass_upper", ctypes.c_int32), ("reserved", ctypes.c_int32)]


class P3B200ConservationFields(ctypes.Structure):
    _fields_ = ([(n, ctypes.c_void_p) for n in CONSERVATION_FIELDS] +
                [("col_stride", ctypes.c_int64), ("wind_col_stride", ctypes.c_int64), ("wind_comp_stride", ctypes.c_int64)])


class P3B200ConservationResult(ctypes.Structure):
    _fields_ = [("mass_rel_err", ctypes.c_double), ("energy_rel_err", ctypes.c_double), ("mass_col", ctypes.c_int64),
                ("energy_col", ctypes.c_int64), ("pass_", ctypes.c_int32), ("mass_below_tol", ctypes.c_int32),
                ("energy_below_tol", ctypes.c_int32), ("reserved", ctypes.c_int32)]


def _check_results(arr):
    keys = ("lb", "ub", "min_val", "max_val", "min_loc", "max_loc", "result", "pass_lower", "pass_upper")
    return {name: {k: getattr(arr[i], k) for k in keys} for i, name in enumerate(CHECK_FIELDS)}


def make_runtime(overrides=None):
    r = P3B200Runtime()
    lib().p3b200_runtime_defaults(ctypes.byref(r))
    for k, v in (overrides or {}).items():
        if k == "do_ice_production":
            r.do_ice_production = int(bool(v))
        else:
            setattr(r, k, float(v))
    return r


def _levels(name, nlev):
    return 1 if name in SURFACE_1D else (nlev + 1 if name in INTERFACE_FIELDS else nlev)


class P3Handle:
    """Owns one p3b200 handle (device tables + diagnostics) on one GPU."""

    def __init__(self, device=0, real_bytes=8, tables=None, flavour="product"):
        self._L = lib(flavour)
        self.flavour = flavour
        self._h = ctypes.c_void_p()
        self.real_bytes = real_bytes
        self.device = device
        t = tables if tables is not None else load_tables()
        check(self._L.p3b200_create(ctypes.byref(t), device, real_bytes, ctypes.byref(self._h)), "p3b200_create")

    def close(self):
        if self._h:
            self._L.p3b200_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- argument marshalling ------------------------------------------------------------------
    def _infra(self, ncol, nlev, dt, predict_nc, prescribed_ccn, zero_flux_top, it=1, zero_history=False):
        return P3B200Infra(float(dt), it, ncol, nlev, int(predict_nc), int(prescribed_ccn), int(zero_flux_top),
                           int(zero_history))

    def fields_struct(self, fields, host):
        """dict name -> (torch cuda tensor | numpy array), level index contiguous."""
        f = P3B200Fields()
        qc = fields["qc"]
        ncol, nlev = qc.shape
        for i, n in enumerate(FIELDS):
            a = fields.get(n)
            if a is None:
                f.ptr[i] = None
                f.col_stride[i] = 0
                continue
            if host:
                assert a.dtype.itemsize == self.real_bytes, n
                f.ptr[i] = a.ctypes.data
                f.col_stride[i] = (a.strides[0] // self.real_bytes) if a.ndim == 2 else 1
                if a.ndim == 2:
                    assert a.strides[1] == self.real_bytes and a.shape == (ncol, _levels(n, nlev)), n
            else:
                assert a.is_cuda and a.element_size() == self.real_bytes, n
                f.ptr[i] = a.data_ptr()
                f.col_stride[i] = a.stride(0) if a.dim() == 2 else 1
                if a.dim() == 2:
                    assert a.stride(1) == 1 and tuple(a.shape) == (ncol, _levels(n, nlev)), n
        return f, ncol, nlev

    # ---- entry points --------------------------------------------------------------------------
    def main(self, fields, dt, predict_nc=True, prescribed_ccn=True, runtime=None, zero_flux_top=True,
             stream=None, timed=False, diag=None, zero_history=False):
        """p3_main on device tensors, in place.  Returns device microseconds if timed else None."""
        f, ncol, nlev = self.fields_struct(fields, host=False)
        infra = self._infra(ncol, nlev, dt, predict_nc, prescribed_ccn, zero_flux_top, zero_history=zero_history)
        us = ctypes.c_int64(0)
        sptr = ctypes.c_void_p(stream) if stream else None
        dptr = ctypes.c_void_p(diag.data_ptr()) if diag is not None else None
        check(self._L.p3b200_main(self._h, ctypes.byref(runtime) if runtime is not None else None, ctypes.byref(f),
                                ctypes.byref(infra), sptr, ctypes.byref(us) if timed else None, dptr), "p3b200_main")
        return us.value if timed else None

    def main_preamble(self, fields, pre_fields, dt, predict_nc=True, prescribed_ccn=True, runtime=None,
                      zero_flux_top=True, stream=None, zero_history=False):
        """p3_preamble fused into p3_main's classification pass (p3b200_main_preamble); `pre_fields` aliases `fields`
        for every array the preamble writes."""
        f, ncol, nlev = self.fields_struct(fields, host=False)
        infra = self._infra(ncol, nlev, dt, predict_nc, prescribed_ccn, zero_flux_top, zero_history=zero_history)
        pf = P3B200PreambleFields()
        for n in PREAMBLE_FIELDS:
            setattr(pf, n, pre_fields[n].data_ptr())
        pf.col_stride = pre_fields["qv"].stride(0)
        sptr = ctypes.c_void_p(stream) if stream else None
        check(self._L.p3b200_main_preamble(self._h, ctypes.byref(runtime) if runtime is not None else None,
                                           ctypes.byref(f), ctypes.byref(infra), ctypes.byref(pf), sptr, None, None),
              "p3b200_main_preamble")

    def main_postamble(self, fields, post_fields, dt, predict_nc=True, prescribed_ccn=True, runtime=None,
                       zero_flux_top=True, stream=None, checks=False):
        """p3_main with p3_postamble fused into the tail of phase C (p3b200_main_postamble).  `post_fields` is a
        postamble dict whose th_atm / mixing ratios / radii / surface fluxes ARE the tensors of `fields`.
        Returns the 15 check results when checks=True."""
        f, ncol, nlev = self.fields_struct(fields, host=False)
        infra = self._infra(ncol, nlev, dt, predict_nc, prescribed_ccn, zero_flux_top)
        pf = self._postamble_struct(post_fields, dt)
        out = (P3B200CheckResult * len(CHECK_FIELDS))() if checks else None
        check(self._L.p3b200_main_postamble(self._h, ctypes.byref(runtime) if runtime is not None else None,
                                          ctypes.byref(f), ctypes.byref(infra), ctypes.byref(pf), out,
                                          ctypes.c_void_p(stream) if stream else None, None, None),
              "p3b200_main_postamble")
        return _check_results(out) if checks else None

    def main_prepared(self, f, infra, stream=None, runtime=None):
        """Hot loop variant for benchmarks: pre-marshalled structs, asynchronous."""
        check(self._L.p3b200_main(self._h, ctypes.byref(runtime) if runtime is not None else None, ctypes.byref(f),
                                ctypes.byref(infra), ctypes.c_void_p(stream) if stream else None, None, None),
              "p3b200_main")

    def main_host(self, fields, dt, predict_nc=True, prescribed_ccn=True, runtime=None, zero_flux_top=True,
                  want_diag=True, zero_history=False):
        """p3_main on host numpy arrays through the chunked H2D/kernel/D2H pipeline, in place.  zero_history: the
        call zero-fills the three history arrays itself (they are then not copied to the device)."""
        f, ncol, nlev = self.fields_struct(fields, host=True)
        infra = self._infra(ncol, nlev, dt, predict_nc, prescribed_ccn, zero_flux_top, zero_history=zero_history)
        us = ctypes.c_int64(0)
        diag = np.full((ncol, NUM_DIAG), -7, dtype=np.int32) if want_diag else None
        check(self._L.p3b200_main_host(self._h, ctypes.byref(runtime) if runtime is not None else None, ctypes.byref(f),
                                     ctypes.byref(infra), ctypes.byref(us),
                                     diag.ctypes.data if want_diag else None), "p3b200_main_host")
        return us.value, diag

    def water_energy_totals(self, fields, dt, stream=None):
        """Domain totals [water, energy, surface water out, surface energy out] of a dict of device tensors
        (p3b200_water_energy_totals); returns a numpy array of 4 doubles."""
        f, ncol, nlev = self.fields_struct(fields, host=False)
        infra = self._infra(ncol, nlev, dt, True, True, True)
        out = np.zeros(4, dtype=np.float64)
        check(self._L.p3b200_water_energy_totals(self._h, ctypes.byref(f), ctypes.byref(infra),
                                               ctypes.c_void_p(stream) if stream else None, None, out.ctypes.data),
              "p3b200_water_energy_totals")
        return out

    def water_energy_totals_static(self, fields, dt, totals4_dev, stream=None):
        """The same totals through the allocation-free entry point (p3b200_water_energy_totals_static): the scratch is
        a fixed device array of the library, the 4 doubles land in `totals4_dev` (a device tensor of the caller's) in
        stream order; nothing is copied back or synchronised.  At most STATIC_BUDGET_MAX_COLS columns."""
        f, ncol, nlev = self.fields_struct(fields, host=False)
        infra = self._infra(ncol, nlev, dt, True, True, True)
        assert totals4_dev.is_cuda and totals4_dev.dtype.itemsize == 8 and totals4_dev.numel() >= 4
        check(self._L.p3b200_water_energy_totals_static(self._h, ctypes.byref(f), ctypes.byref(infra),
                                                      ctypes.c_void_p(stream) if stream else None,
                                                      ctypes.c_void_p(totals4_dev.data_ptr())),
              "p3b200_water_energy_totals_static")

    def preamble(self, fields, stream=None):
        """p3_preamble on device tensors (dict keyed by PREAMBLE_FIELDS, all (ncol, nlev) with one column stride)."""
        f = P3B200PreambleFields()
        ncol, nlev = fields["qv"].shape
        for n in PREAMBLE_FIELDS:
            t = fields[n]
            assert t.is_cuda and t.stride(1) == 1 and t.stride(0) == fields["qv"].stride(0), n
            setattr(f, n, t.data_ptr())
        f.col_stride = fields["qv"].stride(0)
        check(self._L.p3b200_preamble(self._h, ctypes.byref(f), ncol, nlev, ctypes.c_void_p(stream or 0)), "p3b200_preamble")

    def postamble(self, fields, dt, stream=None):
        """p3_postamble on device tensors (POSTAMBLE_FIELDS_2D + POSTAMBLE_FIELDS_1D; the four boundary fluxes may be absent)."""
        f = self._postamble_struct(fields, dt)
        ncol, nlev = fields["qv"].shape
        check(self._L.p3b200_postamble(self._h, ctypes.byref(f), ncol, nlev, ctypes.c_void_p(stream or 0)), "p3b200_postamble")

    @staticmethod
    def _postamble_struct(fields, dt):
        f = P3B200PostambleFields()
        ncol, nlev = fields["qv"].shape
        for n in POSTAMBLE_FIELDS_2D:
            t = fields[n]
            assert t.is_cuda and t.stride(1) == 1 and t.stride(0) == fields["qv"].stride(0), n
            setattr(f, n, t.data_ptr())
        for n in POSTAMBLE_FIELDS_1D:
            setattr(f, n, fields[n].data_ptr() if n in fields and fields[n] is not None else None)
        f.col_stride = fields["qv"].stride(0)
        f.dt = float(dt)
        return f

    def postamble_checked(self, fields, dt, stream=None):
        """p3_postamble with the 15 property checks of the P3 process folded into the same pass.
        Returns {check name: {lb, ub, min_val, max_val, min_loc, max_loc, result, pass_lower, pass_upper}}."""
        f = self._postamble_struct(fields, dt)
        ncol, nlev = fields["qv"].shape
        out = (P3B200CheckResult * len(CHECK_FIELDS))()
        check(self._L.p3b200_postamble_checked(self._h, ctypes.byref(f), ncol, nlev, out, ctypes.c_void_p(stream or 0)),
              "p3b200_postamble_checked")
        return _check_results(out)

    def property_checks(self, fields, stream=None):
        """The 15 property checks as one stand-alone pass over device tensors named like CHECK_FIELDS."""
        f = P3B200CheckFields()
        ncol, nlev = fields["qv"].shape
        for n in CHECK_FIELDS:
            t = fields[n]
            assert t.is_cuda and (n in CHECK_COLUMN_FIELDS or (t.stride(1) == 1 and t.stride(0) == fields["qv"].stride(0))), n
            setattr(f, n, t.data_ptr())
        f.col_stride = fields["qv"].stride(0)
        out = (P3B200CheckResult * len(CHECK_FIELDS))()
        check(self._L.p3b200_property_checks(self._h, ctypes.byref(f), ncol, nlev, out, ctypes.c_void_p(stream or 0)),
              "p3b200_property_checks")
        return _check_results(out)

    def repair_qv(self, qv, stream=None):
        ncol, nlev = qv.shape
        check(self._L.p3b200_repair_qv(self._h, ctypes.c_void_p(qv.data_ptr()), ctypes.c_int64(qv.stride(0)), ncol, nlev,
                                     ctypes.c_void_p(stream or 0)), "p3b200_repair_qv")

    def _conservation_struct(self, fields):
        f = P3B200ConservationFields()
        ncol, nlev = fields["qv"].shape
        for n in CONSERVATION_FIELDS:
            t = fields.get(n)
            setattr(f, n, t.data_ptr() if t is not None else None)
        f.col_stride = fields["qv"].stride(0)
        w = fields["horiz_winds"]
        f.wind_col_stride, f.wind_comp_stride = w.stride(0), w.stride(1)
        return f, ncol, nlev

    def column_conservation_begin(self, fields, stream=None):
        """compute_current_mass / compute_current_energy of MassAndEnergyColumnConservationCheck on device tensors."""
        f, ncol, nlev = self._conservation_struct(fields)
        check(self._L.p3b200_column_conservation_begin(self._h, ctypes.byref(f), ncol, nlev,
                                                       ctypes.c_void_p(stream) if stream else None), "conservation_begin")

    def column_conservation_check(self, fields, dt, mass_tol, energy_tol, stream=None):
        f, ncol, nlev = self._conservation_struct(fields)
        r = P3B200ConservationResult()
        check(self._L.p3b200_column_conservation_check(self._h, ctypes.byref(f), ncol, nlev, ctypes.c_double(dt),
                                                       ctypes.c_double(mass_tol), ctypes.c_double(energy_tol),
                                                       ctypes.byref(r), ctypes.c_void_p(stream) if stream else None),
              "conservation_check")
        out = {k: getattr(r, k) for k, _ in r._fields_}
        arrs = [np.zeros(ncol) for _ in range(4)]
        check(self._L.p3b200_column_conservation_get(self._h, ncol, *[a.ctypes.data_as(ctypes.c_void_p) for a in arrs]),
              "conservation_get")
        out["mass_prev"], out["energy_prev"], out["rel_mass"], out["rel_energy"] = arrs
        return out

    def get_diag(self, ncol):
        out = np.empty((ncol, NUM_DIAG), dtype=np.int32)
        check(self._L.p3b200_get_diag(self._h, out.ctypes.data, ncol), "p3b200_get_diag")
        return out

    def error_summary(self):
        v = ctypes.c_int32(0)
        check(self._L.p3b200_error_summary(self._h, ctypes.byref(v)), "p3b200_error_summary")
        return v.value

    def last_kernel_ms(self):
        ms = ctypes.c_double(0)
        n = ctypes.c_int64(0)
        check(self._L.p3b200_last_kernel_ms(self._h, ctypes.byref(ms), ctypes.byref(n)), "p3b200_last_kernel_ms")
        return ms.value, n.value

    def last_phase_ms(self):
        """Device ms of the phases of the last call: phase A, cloud/rain/ice sedimentation, phase C, and the whole
        sedimentation stage (the species run side by side unless P3B200_SED_SERIAL=1, see include/p3_b200.h)."""
        ms = (ctypes.c_double * 6)()
        check(self._L.p3b200_last_phase_ms(self._h, ms), "p3b200_last_phase_ms")
        out = dict(zip(("phase_a", "sed_cloud", "sed_rain", "sed_ice", "phase_c", "sed_stage"), [float(x) for x in ms]))
        out["sed_concurrent"] = int(self._L.p3b200_last_sed_concurrent(self._h))
        a3 = (ctypes.c_double * 3)()
        check(self._L.p3b200_last_phase_a_ms(self._h, a3), "p3b200_last_phase_a_ms")
        out["a_classify"], out["a_part2"], out["a_finish"] = (float(x) for x in a3)
        return out


def to_device(fields, device="cuda:0"):
    import torch
    return {n: torch.from_numpy(np.ascontiguousarray(v)).to(device) for n, v in fields.items()}


def to_host(fields):
    return {n: v.detach().cpu().numpy() for n, v in fields.items()}
