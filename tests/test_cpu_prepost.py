"""p3_preamble / p3_postamble restatement (oracle/p3_oracle_prepost.cpp) against closed forms written here with
numpy from the reference formulas (eamxx_p3_process_interface.hpp:62-262), plus the wet->dry->wet round trip."""
import numpy as np
import pytest

from scream_b200 import prepost_cases
from scream_b200.fields import POSTAMBLE_OUTPUTS, PREAMBLE_OUTPUTS


@pytest.mark.parametrize("dtype,tol", [(np.float64, 4e-16), (np.float32, 3e-7)])
def test_preamble_closed_forms(oracle, dtype, tol):
    f = prepost_cases.make_preamble(37, 72, dtype)
    g = {k: v.copy() for k, v in f.items()}
    oracle.preamble(g)
    pd, pdd, pmid, T = (f[n].astype(np.float64) for n in ("pseudo_density", "pseudo_density_dry", "pmid", "T_atm"))
    for n in ("qv", "qc", "nc", "qr", "nr", "qi", "qm", "ni", "bm", "qv_prev"):
        np.testing.assert_allclose(g[n], f[n].astype(np.float64) * pd / pdd, rtol=4 * tol, atol=0)
    exner = (pmid / 1.0e5) ** (287.042 / 1004.64)
    np.testing.assert_allclose(g["inv_exner"], 1 / exner, rtol=1e-6 if dtype == np.float32 else 1e-14)
    np.testing.assert_allclose(g["th_atm"], T / exner, rtol=1e-6 if dtype == np.float32 else 1e-14)
    tv = T * (1 + (-1 + 28.966 / 18.016) * f["qv"].astype(np.float64))
    np.testing.assert_allclose(g["dz"], 287.042 / 9.80616 * pd * tv / pmid, rtol=1e-6 if dtype == np.float32 else 1e-14)
    cf = np.maximum(f["cld_frac_t"], dtype(0.0001))
    assert np.array_equal(g["cld_frac_l"], cf) and np.array_equal(g["cld_frac_i"], cf)
    cfr = cf.copy()
    cfr[:, 1:] = np.maximum(cf[:, 1:], f["cld_frac_t"][:, :-1])
    assert np.array_equal(g["cld_frac_r"], cfr)
    for n in ("pmid", "pseudo_density", "pseudo_density_dry", "T_atm", "cld_frac_t"):
        assert np.array_equal(g[n], f[n]), n      # inputs untouched
    assert set(PREAMBLE_OUTPUTS) <= set(g)


@pytest.mark.parametrize("with_fluxes", [True, False])
def test_postamble_closed_forms_and_round_trip(oracle, with_fluxes):
    f = prepost_cases.make_postamble(29, 128, np.float64, with_fluxes=with_fluxes)
    g = {k: v.copy() for k, v in f.items()}
    oracle.postamble(g, 300.0)
    pd, pdd, pmid = f["pseudo_density"], f["pseudo_density_dry"], f["pmid"]
    exner = (pmid / 1.0e5) ** (287.042 / 1004.64)
    T = (f["th_atm"] * exner - f["T_atm"]) * pdd / pd + f["T_atm"]
    np.testing.assert_allclose(g["T_atm"], T, rtol=1e-14)
    assert np.array_equal(g["T_prev"], g["T_atm"]) and np.array_equal(g["qv_prev"], g["qv"])
    for n in ("qv", "qc", "nc", "qr", "nr", "qi", "qm", "ni", "bm"):
        np.testing.assert_allclose(g[n], f[n] * pdd / pd, rtol=1e-15, atol=0)
    for n in ("diag_eff_radius_qc", "diag_eff_radius_qi", "diag_eff_radius_qr"):
        np.testing.assert_allclose(g[n], f[n] * 1e6, rtol=1e-15)
    np.testing.assert_allclose(g["precip_liq_surf_mass"], f["precip_liq_surf_mass"] + f["precip_liq_surf_flux"] * 1000.0 * 300.0, rtol=1e-15)
    np.testing.assert_allclose(g["precip_ice_surf_mass"], f["precip_ice_surf_mass"] + f["precip_ice_surf_flux"] * 1000.0 * 300.0, rtol=1e-15)
    if with_fluxes:
        assert np.all(g["vapor_flux"] == 0) and np.all(g["heat_flux"] == 0)
        assert np.array_equal(g["water_flux"], f["precip_liq_surf_flux"] + f["precip_ice_surf_flux"])
        assert np.array_equal(g["ice_flux"], f["precip_ice_surf_flux"])
    assert set(n for n in POSTAMBLE_OUTPUTS if with_fluxes or not n.endswith("_flux")) <= set(g)

    # preamble(postamble(x)) restores the dry mixing ratios to rounding (wet -> dry is the inverse map)
    pre = prepost_cases.make_preamble(29, 128, np.float64, seed=20241105)
    h = {k: v.copy() for k, v in pre.items()}
    for n in ("qv", "qc", "nc", "qr", "nr", "qi", "qm", "ni", "bm"):
        h[n] = g[n].copy()
    oracle.preamble(h)
    for n in ("qc", "nc", "qr", "nr", "qi", "qm", "ni", "bm"):
        np.testing.assert_allclose(h[n], f[n], rtol=1e-15, atol=0)
