"""The reference's own C bridge, p3_main_f (p3_functions_f90.hpp:945-952; implementation p3_functions_f90.cpp:1888-2103),
as re-implemented by p3b200_p3_main_f: column-major ("Fortran") host arrays, 1-based its/kts, bool flags, outputs
zeroed on entry, P3Runtime defaults, elapsed microseconds as return value.  Called here the way the reference's
p3_run_and_cmp / run_bfb tests call theirs, on the reference's test case ic::Factory::mixed(3, 72), and compared with the
reference's own outputs (tests/golden/ref_mixed_3x72_dt300_6steps.npz, generated from oracle/_ref)."""
import ctypes
import os

import numpy as np
import pytest

from scream_b200 import ic_cases
from scream_b200.fields import OUTPUTS
from tests import parity

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_mixed_3x72_dt300_6steps.npz")
FLAGS = [(0, 0), (0, 1), (1, 0), (1, 1)]

# argument order of p3_main_f; None marks the scalar slots
ORDER = ["qc", "nc", "qr", "nr", "th", "qv", "<dt>", "qi", "qm", "ni", "bm", "pres", "dz", "nc_nuceat_tend", "nccn",
         "ni_activated", "inv_qc_relvar", "<it>", "precip_liq_surf", "precip_ice_surf", "<its>", "<ite>", "<kts>", "<kte>",
         "diag_eff_radius_qc", "diag_eff_radius_qi", "diag_eff_radius_qr", "rho_qi", "<predict_nc>", "<prescribed_ccn>",
         "dpres", "inv_exner", "qv2qi_depos_tend", "precip_liq_flux", "precip_ice_flux", "cld_frac_r", "cld_frac_l",
         "cld_frac_i", "liq_ice_exchange", "vap_liq_exchange", "vap_ice_exchange", "qv_prev", "t_prev"]
NOT_IN_BRIDGE = ("precip_total_tend", "nevapr")     # internal to p3_main_f (p3_functions_f90.cpp:1972-1975)


def call_bridge(L, f, dt, pn, pc, it=1):
    ncol, nlev = f["qc"].shape
    colmajor = {n: np.asfortranarray(f[n]) for n in ORDER if not n.startswith("<")}
    garbage = ("qv2qi_depos_tend", "diag_eff_radius_qc", "diag_eff_radius_qi", "diag_eff_radius_qr", "rho_qi",
               "precip_liq_flux", "precip_ice_flux", "liq_ice_exchange", "vap_liq_exchange", "vap_ice_exchange",
               "precip_liq_surf", "precip_ice_surf")
    for n in garbage:          # the bridge zeroes its outputs on entry: whatever the caller left must not matter
        colmajor[n][...] = 777.0
    scal = {"<dt>": ctypes.c_double(dt), "<it>": it, "<its>": 1, "<ite>": ncol, "<kts>": 1, "<kte>": nlev,
            "<predict_nc>": bool(pn), "<prescribed_ccn>": bool(pc)}
    args = [scal[n] if n.startswith("<") else colmajor[n].ctypes.data_as(ctypes.POINTER(ctypes.c_double)) for n in ORDER]
    rc = L.p3b200_p3_main_f(*args)
    return rc, {n: np.ascontiguousarray(v) for n, v in colmajor.items()}


@pytest.fixture(scope="module")
def bridge(p3lib):
    L = p3lib.lib()
    t = p3lib.load_tables()
    assert L.p3b200_set_default_tables(ctypes.byref(t), 0) == 0
    yield L
    L.p3b200_set_default_tables(None, 0)


@pytest.mark.parametrize("pn,pc", FLAGS)
def test_p3_main_f_matches_the_reference_on_its_own_test_case(bridge, pn, pc):
    g = np.load(GOLDEN)
    f = ic_cases.make_mixed(3, 72)
    rc, got = call_bridge(bridge, f, 300.0, pn, pc)
    assert rc >= 0, rc                                  # elapsed microseconds
    names = [n for n in OUTPUTS if n not in NOT_IN_BRIDGE]
    ref = {n: g[f"f64/libm/nc{pn}_ccn{pc}/step0/{n}"] for n in names}
    errs = parity.field_errors(ref, got, names)
    assert max(errs.values()) <= parity.TOL[8], {k: v for k, v in errs.items() if v > parity.TOL[8]}
    for n in ("pres", "dz", "dpres", "inv_exner", "cld_frac_i", "qv_prev"):    # inputs come back untouched
        assert np.array_equal(got[n], f[n]), n


def test_p3_main_f_rejects_non_unit_lower_bounds(bridge):
    f = ic_cases.make_mixed(3, 72)
    ncol, nlev = f["qc"].shape
    colmajor = {n: np.asfortranarray(f[n]) for n in ORDER if not n.startswith("<")}
    scal = {"<dt>": ctypes.c_double(300.0), "<it>": 1, "<its>": 2, "<ite>": ncol, "<kts>": 1, "<kte>": nlev,
            "<predict_nc>": True, "<prescribed_ccn>": True}
    args = [scal[n] if n.startswith("<") else colmajor[n].ctypes.data_as(ctypes.POINTER(ctypes.c_double)) for n in ORDER]
    assert bridge.p3b200_p3_main_f(*args) < 0      # the reference asserts its == kts == 1 (p3_functions_f90.cpp:1911-1912)


def test_p3_main_f_six_steps_tracks_the_reference(bridge):
    g = np.load(GOLDEN)
    f = ic_cases.make_mixed(3, 72)
    names = [n for n in OUTPUTS if n not in NOT_IN_BRIDGE]
    for step in range(6):
        rc, got = call_bridge(bridge, f, 300.0, 1, 1, it=step + 1)
        assert rc >= 0
        ref = {n: g[f"f64/libm/nc1_ccn1/step{step}/{n}"] for n in names}
        errs = parity.field_errors(ref, got, names)
        assert max(errs.values()) <= 1e-9, (step, {k: v for k, v in errs.items() if v > 1e-9})
        for n in got:
            f[n] = got[n]
