"""Synthetic inputs for p3_preamble / p3_postamble (P3Microphysics::run_impl, eamxx_p3_run.cpp): the fields the
functors read from the field manager, seeded and dimensioned like the p3_main cases (level 0 = model top)."""
import numpy as np

from .fields import POSTAMBLE_FIELDS_1D, POSTAMBLE_FIELDS_2D, PREAMBLE_FIELDS


def make_preamble(ncol, nlev, dtype=np.float64, seed=20241104):
    rng = np.random.default_rng(seed)
    k = np.arange(nlev, dtype=np.float64)[None, :]
    f = {}
    f["pmid"] = 100.0 + 1.0e5 * (k + 0.5) / nlev + rng.uniform(-20, 20, (ncol, nlev))
    f["pseudo_density"] = np.full((ncol, nlev), 1.0e5 / nlev) * rng.uniform(0.9, 1.1, (ncol, nlev))
    f["T_atm"] = 190.0 + 110.0 * k / max(nlev - 1, 1) + rng.uniform(-3, 3, (ncol, nlev))
    f["cld_frac_t"] = np.where(rng.random((ncol, nlev)) < 0.5, rng.uniform(0, 1, (ncol, nlev)), 0.0)
    f["qv"] = rng.uniform(1e-6, 2e-2, (ncol, nlev))
    f["pseudo_density_dry"] = f["pseudo_density"] * (1.0 - f["qv"] - rng.uniform(0, 1e-3, (ncol, nlev)))
    for n, hi in (("qc", 1e-3), ("qr", 1e-4), ("qi", 1e-3), ("qm", 1e-3), ("bm", 1e-5)):
        f[n] = np.where(rng.random((ncol, nlev)) < 0.4, rng.uniform(0, hi, (ncol, nlev)), 0.0)
    for n in ("nc", "nr", "ni"):
        f[n] = rng.uniform(0, 1e8, (ncol, nlev))
    f["qv_prev"] = f["qv"] * rng.uniform(0.98, 1.02, (ncol, nlev))
    for n in ("inv_exner", "th_atm", "cld_frac_l", "cld_frac_i", "cld_frac_r", "dz"):
        f[n] = np.full((ncol, nlev), -1.0)        # outputs: poisoned
    return {n: np.ascontiguousarray(f[n], dtype=dtype) for n in PREAMBLE_FIELDS}


def make_postamble(ncol, nlev, dtype=np.float64, seed=20241105, with_fluxes=True):
    pre = make_preamble(ncol, nlev, np.float64, seed)
    rng = np.random.default_rng(seed + 1)
    f = {n: pre[n] for n in ("pmid", "pseudo_density", "pseudo_density_dry", "T_atm", "qv", "qc", "nc", "qr", "nr",
                             "qi", "qm", "ni", "bm")}
    exner = (f["pmid"] / 1.0e5) ** (287.042 / 1004.64)
    f["th_atm"] = (f["T_atm"] + rng.uniform(-2, 2, (ncol, nlev))) / exner      # theta after p3_main
    f["diag_eff_radius_qc"] = rng.uniform(1e-6, 5e-5, (ncol, nlev))
    f["diag_eff_radius_qi"] = rng.uniform(1e-6, 5e-4, (ncol, nlev))
    f["diag_eff_radius_qr"] = rng.uniform(1e-5, 5e-3, (ncol, nlev))
    f["T_prev"] = np.full((ncol, nlev), -1.0)
    f["qv_prev"] = np.full((ncol, nlev), -1.0)
    f["precip_liq_surf_flux"] = rng.uniform(0, 1e-6, ncol)
    f["precip_ice_surf_flux"] = rng.uniform(0, 1e-7, ncol)
    f["precip_liq_surf_mass"] = rng.uniform(0, 1e-2, ncol)
    f["precip_ice_surf_mass"] = rng.uniform(0, 1e-3, ncol)
    if with_fluxes:
        for n in ("vapor_flux", "water_flux", "ice_flux", "heat_flux"):
            f[n] = np.full(ncol, -1.0)
    names = POSTAMBLE_FIELDS_2D + POSTAMBLE_FIELDS_1D
    return {n: np.ascontiguousarray(f[n], dtype=dtype) for n in names if n in f}
