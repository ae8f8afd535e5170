"""MassAndEnergyColumnConservationCheck on the GPU (p3b200_column_conservation_begin / _check, p3_budget.cu) against the
numpy restatement (oracle/p3_oracle_checks.py): per-column before-values and relative errors bit for bit (the kernel
adds the levels in index order like the reference's serialised reduction), MaxLoc value and column, tolerance logic;
and wrapped around a real preamble -> p3_main -> postamble step, where the boundary fluxes of the postamble close the
budget."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import p3_oracle_checks as oc  # noqa: E402

pytestmark = pytest.mark.gpu


def make_state(ncol, nlev, dtype, seed=1):
    rng = np.random.default_rng(seed)
    f = {"pseudo_density": rng.uniform(300, 1200, (ncol, nlev)), "T_mid": rng.uniform(190, 305, (ncol, nlev)),
         "horiz_winds": rng.uniform(-40, 40, (ncol, 2, nlev))}
    for n, hi in (("qv", 2e-2), ("qc", 1e-3), ("qi", 1e-3), ("qr", 1e-4)):
        f[n] = rng.uniform(0, hi, (ncol, nlev))
    f["ps"] = rng.uniform(9e4, 1.05e5, ncol); f["phis"] = rng.uniform(0, 3e4, ncol)
    for n in ("vapor_flux", "water_flux", "ice_flux", "heat_flux"):
        f[n] = np.zeros(ncol)
    return {k: np.ascontiguousarray(v, dtype=dtype) for k, v in f.items()}


def to_dev(f):
    import torch
    return {k: torch.from_numpy(v).cuda() for k, v in f.items()}


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
@pytest.mark.parametrize("ncol,nlev", [(257, 72), (1000, 128), (5, 37)])
def test_column_check_matches_the_restatement(p3lib, dtype, ncol, nlev):
    rb = np.dtype(dtype).itemsize
    h = p3lib.P3Handle(0, rb)
    f = make_state(ncol, nlev, dtype)
    h.column_conservation_begin(to_dev(f))
    m0, e0 = oc.column_mass(f), oc.column_energy(f)
    # the process: move water around, let some leave through the surface, with and without reporting it
    g = {k: v.copy() for k, v in f.items()}
    rng = np.random.default_rng(2)
    g["qr"] *= dtype(0.5)
    g["qv"] += (rng.uniform(0, 1e-5, g["qv"].shape)).astype(dtype)
    g["T_mid"] += rng.uniform(-0.5, 0.5, g["T_mid"].shape).astype(dtype)
    g["water_flux"] = rng.uniform(0, 1e-7, ncol).astype(dtype)
    g["vapor_flux"] = rng.uniform(0, 1e-5, ncol).astype(dtype)
    g["ice_flux"] = rng.uniform(0, 1e-8, ncol).astype(dtype)
    g["heat_flux"] = rng.uniform(-10, 10, ncol).astype(dtype)
    g["qv"][ncol // 2, 3] = np.nan                      # a NaN column must not win the MaxLoc
    dt = 300.0
    got = h.column_conservation_check(to_dev(g), dt, 1e-3, 1e-3)
    ref = oc.column_conservation_check(g, m0, e0, dt, 1e-3, 1e-3)
    assert np.array_equal(got["mass_prev"].astype(dtype), m0) and np.array_equal(got["energy_prev"].astype(dtype), e0)
    assert np.array_equal(got["rel_mass"].astype(dtype), ref["rel_mass"], equal_nan=True)
    assert np.array_equal(got["rel_energy"].astype(dtype), ref["rel_energy"], equal_nan=True)
    assert got["mass_rel_err"] == ref["mass_rel_err"] and got["mass_col"] == ref["mass_col"]
    assert got["energy_rel_err"] == ref["energy_rel_err"] and got["energy_col"] == ref["energy_col"]
    assert bool(got["pass_"]) == ref["pass_"]
    # unchanged state, zero fluxes: both errors exactly zero, check passes at any positive tolerance
    got = h.column_conservation_check(to_dev(f), dt, 1e-300, 1e-300)
    assert got["pass_"] == 1 and got["mass_rel_err"] == 0.0 and got["energy_rel_err"] == 0.0
    h.close()


def test_check_closes_around_a_run_impl_step(p3lib):
    """compute_current_* -> preamble -> p3_main -> postamble (with boundary fluxes) -> check(): P3 conserves column
    water and energy up to the surface fluxes it reports, so the check passes at the driver's tolerances and fails
    when the reported water flux is withheld."""
    import torch
    from scream_b200 import ic_cases
    from tests import test_gpu_process as tp
    ncol, nlev, dt = 512, 72, 300.0
    fm = tp.field_manager_state(ncol, nlev, seed=9)
    h = p3lib.P3Handle(0, 8)
    d = {k: torch.from_numpy(v).cuda() for k, v in fm.items()}
    cons = {"pseudo_density": d["pseudo_density"], "T_mid": d["T_mid"], "qv": d["qv"], "qc": d["qc"], "qi": d["qi"], "qr": d["qr"],
            "horiz_winds": torch.zeros(ncol, 2, nlev, dtype=torch.float64, device="cuda"),
            "ps": torch.full((ncol,), 1.0e5, dtype=torch.float64, device="cuda"),
            "phis": torch.zeros(ncol, dtype=torch.float64, device="cuda"),
            "vapor_flux": d["vapor_flux"], "water_flux": d["water_flux"], "ice_flux": d["ice_flux"], "heat_flux": d["heat_flux"]}
    h.column_conservation_begin(cons)
    buf = {n: torch.empty(ncol, nlev, dtype=torch.float64, device="cuda") for n in ("inv_exner", "th_atm", "cld_frac_l", "cld_frac_i", "dz")}
    pre = {"pmid": d["p_mid"], "pseudo_density": d["pseudo_density"], "pseudo_density_dry": d["pseudo_density_dry"],
           "T_atm": d["T_mid"], "cld_frac_t": d["cldfrac_tot"], "qv_prev": d["qv_prev_micro_step"], "cld_frac_r": d["rainfrac"], **buf}
    for n in ("qv", "qc", "nc", "qr", "nr", "qi", "qm", "ni", "bm"):
        pre[n] = d[n]
    h.preamble(pre)
    z = lambda *s: torch.zeros(*s, dtype=torch.float64, device="cuda")  # noqa: E731
    main = {"qc": d["qc"], "nc": d["nc"], "qr": d["qr"], "nr": d["nr"], "qi": d["qi"], "qm": d["qm"], "ni": d["ni"], "bm": d["bm"],
            "qv": d["qv"], "th": buf["th_atm"], "nc_nuceat_tend": d["nc_nuceat_tend"], "nccn": d["nccn"],
            "ni_activated": d["ni_activated"], "inv_qc_relvar": d["inv_qc_relvar"], "cld_frac_i": buf["cld_frac_i"],
            "cld_frac_l": buf["cld_frac_l"], "cld_frac_r": d["rainfrac"], "pres": d["p_dry_mid"], "dz": buf["dz"],
            "dpres": d["pseudo_density_dry"], "inv_exner": buf["inv_exner"], "qv_prev": d["qv_prev_micro_step"],
            "t_prev": d["T_prev_micro_step"], "qv2qi_depos_tend": z(ncol, nlev), "diag_eff_radius_qc": d["eff_radius_qc"],
            "diag_eff_radius_qi": d["eff_radius_qi"], "diag_eff_radius_qr": d["eff_radius_qr"], "rho_qi": z(ncol, nlev),
            "precip_liq_flux": z(ncol, nlev + 1), "precip_ice_flux": z(ncol, nlev + 1), "precip_total_tend": d["precip_total_tend"],
            "nevapr": d["nevapr"], "liq_ice_exchange": d["micro_liq_ice_exchange"], "vap_liq_exchange": d["micro_vap_liq_exchange"],
            "vap_ice_exchange": d["micro_vap_ice_exchange"], "precip_liq_surf": z(ncol), "precip_ice_surf": z(ncol)}
    h.main(main, dt, True, True, zero_history=True)
    post = {"pmid": d["p_mid"], "pseudo_density": d["pseudo_density"], "pseudo_density_dry": d["pseudo_density_dry"],
            "th_atm": buf["th_atm"], "T_atm": d["T_mid"], "T_prev": d["T_prev_micro_step"], "qv_prev": d["qv_prev_micro_step"],
            "diag_eff_radius_qc": d["eff_radius_qc"], "diag_eff_radius_qi": d["eff_radius_qi"], "diag_eff_radius_qr": d["eff_radius_qr"],
            "precip_liq_surf_flux": main["precip_liq_surf"], "precip_ice_surf_flux": main["precip_ice_surf"],
            "precip_liq_surf_mass": d["precip_liq_surf_mass"], "precip_ice_surf_mass": d["precip_ice_surf_mass"],
            "vapor_flux": d["vapor_flux"], "water_flux": d["water_flux"], "ice_flux": d["ice_flux"], "heat_flux": d["heat_flux"]}
    for n in ("qv", "qc", "nc", "qr", "nr", "qi", "qm", "ni", "bm"):
        post[n] = d[n]
    h.postamble(post, dt)
    r = h.column_conservation_check(cons, dt, 1e-10, 1e-10)
    print("max relative errors: mass %.3e (col %d), energy %.3e (col %d)" % (r["mass_rel_err"], r["mass_col"], r["energy_rel_err"], r["energy_col"]))
    assert r["pass_"] == 1, r
    if float(d["water_flux"].abs().max()) > 0:
        d["water_flux"].zero_()
        r = h.column_conservation_check(cons, dt, 1e-10, 1e-10)
        assert r["pass_"] == 0
    h.close()
