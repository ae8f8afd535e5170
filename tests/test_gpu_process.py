"""The process-level face scream::P3Microphysics (include/p3_b200_process.hpp, driven by tests/cpp/p3_process_b200.cpp)
against the oracle's restatement of P3Microphysics::run_impl (eamxx_p3_run.cpp:5-46): p3_preamble -> zero the history
fields -> p3_main -> p3_postamble, then the 15 property checks of initialize_impl
(eamxx_p3_process_interface.cpp:226-243) with the repair of qv."""
import os
import subprocess
import sys

import numpy as np
import pytest

from scream_b200 import ic_cases, lib as p3
from scream_b200.fields import CHECKS, FIELDS
from tests import parity

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import p3_oracle_checks as oc  # noqa: E402

EXE = os.path.join(ROOT, "build", "p3_process_b200")
NAMES2D = ["p_mid", "p_dry_mid", "pseudo_density", "pseudo_density_dry", "cldfrac_tot", "nc_nuceat_tend", "nccn",
           "ni_activated", "inv_qc_relvar", "T_mid", "qv", "qc", "qr", "qi", "qm", "nc", "nr", "ni", "bm",
           "qv_prev_micro_step", "T_prev_micro_step", "eff_radius_qc", "eff_radius_qi", "eff_radius_qr",
           "precip_total_tend", "nevapr", "rainfrac", "micro_liq_ice_exchange", "micro_vap_liq_exchange",
           "micro_vap_ice_exchange"]
NAMES1D = ["precip_liq_surf_mass", "precip_ice_surf_mass", "vapor_flux", "water_flux", "ice_flux", "heat_flux"]


def _build():
    subprocess.check_call(["make", "-C", os.path.join(ROOT, "tests", "cpp"), "-s"])


def field_manager_state(ncol, nlev, seed=5, ic="random"):
    """Wet mixing ratios and thermodynamic state as the field manager holds them before run_impl."""
    base = ic_cases.make_random(ncol, nlev, seed=seed) if ic == "random" else ic_cases.make_mixed(ncol, nlev)
    rng = np.random.default_rng(seed)
    fm = {}
    fm["p_mid"] = base["pres"].copy()
    fm["p_dry_mid"] = base["pres"] * (1 - 0.6 * base["qv"])
    fm["pseudo_density"] = base["dpres"].copy()
    fm["pseudo_density_dry"] = base["dpres"] * (1 - base["qv"])
    fm["cldfrac_tot"] = np.where(rng.random((ncol, nlev)) < 0.5, rng.uniform(0, 1, (ncol, nlev)), 0.0)
    for n in ("nc_nuceat_tend", "nccn", "ni_activated", "inv_qc_relvar"):
        fm[n] = base[n].copy()
    fm["T_mid"] = base["th"] / base["inv_exner"]
    for n in ("qv", "qc", "qr", "qi", "qm", "nc", "nr", "ni", "bm"):
        fm[n] = base[n].copy()
    fm["qv_prev_micro_step"] = base["qv_prev"].copy()
    fm["T_prev_micro_step"] = base["t_prev"].copy()
    for n in ("eff_radius_qc", "eff_radius_qi", "eff_radius_qr", "precip_total_tend", "nevapr", "rainfrac"):
        fm[n] = np.full((ncol, nlev), -5.0)
    for n in ("micro_liq_ice_exchange", "micro_vap_liq_exchange", "micro_vap_ice_exchange"):
        fm[n] = np.full((ncol, nlev), 9.0)           # run_impl zero-fills them
    fm["precip_liq_surf_mass"] = rng.uniform(0, 1e-2, ncol)
    fm["precip_ice_surf_mass"] = rng.uniform(0, 1e-3, ncol)
    for n in ("vapor_flux", "water_flux", "ice_flux", "heat_flux"):
        fm[n] = np.full(ncol, -1.0)
    return {k: np.ascontiguousarray(v, dtype=np.float64) for k, v in fm.items()}


def oracle_run_impl(oracle, fm, dt, predict_nc, prescribed_ccn, fluxes):
    """One run_impl on the field-manager dict, in place, through the oracle's three restatements."""
    ncol, nlev = fm["qv"].shape
    pre = {"pmid": fm["p_mid"], "pseudo_density": fm["pseudo_density"], "pseudo_density_dry": fm["pseudo_density_dry"],
           "T_atm": fm["T_mid"], "cld_frac_t": fm["cldfrac_tot"], "qv_prev": fm["qv_prev_micro_step"],
           "cld_frac_r": fm["rainfrac"]}
    for n in ("qv", "qc", "nc", "qr", "nr", "qi", "qm", "ni", "bm"):
        pre[n] = fm[n]
    for n in ("inv_exner", "th_atm", "cld_frac_l", "cld_frac_i", "dz"):
        pre[n] = np.full((ncol, nlev), np.nan)
    oracle.preamble(pre)
    main = {n: None for n in FIELDS}
    for n in ("qc", "nc", "qr", "nr", "qi", "qm", "ni", "bm", "qv"):
        main[n] = fm[n]
    main["th"] = pre["th_atm"]
    for n in ("nc_nuceat_tend", "nccn", "ni_activated", "inv_qc_relvar"):
        main[n] = fm[n]
    main["cld_frac_i"], main["cld_frac_l"], main["cld_frac_r"] = pre["cld_frac_i"], pre["cld_frac_l"], fm["rainfrac"]
    main["pres"], main["dz"], main["dpres"], main["inv_exner"] = fm["p_dry_mid"], pre["dz"], fm["pseudo_density_dry"], pre["inv_exner"]
    main["qv_prev"], main["t_prev"] = fm["qv_prev_micro_step"], fm["T_prev_micro_step"]
    main["qv2qi_depos_tend"] = np.zeros((ncol, nlev)); main["rho_qi"] = np.zeros((ncol, nlev))
    main["diag_eff_radius_qc"], main["diag_eff_radius_qi"], main["diag_eff_radius_qr"] = (fm["eff_radius_qc"], fm["eff_radius_qi"],
                                                                                        fm["eff_radius_qr"])
    main["precip_liq_flux"] = np.zeros((ncol, nlev + 1)); main["precip_ice_flux"] = np.zeros((ncol, nlev + 1))
    main["precip_total_tend"], main["nevapr"] = fm["precip_total_tend"], fm["nevapr"]
    main["liq_ice_exchange"], main["vap_liq_exchange"], main["vap_ice_exchange"] = (fm["micro_liq_ice_exchange"],
                                                                                   fm["micro_vap_liq_exchange"],
                                                                                   fm["micro_vap_ice_exchange"])
    main["precip_liq_surf"] = np.zeros(ncol); main["precip_ice_surf"] = np.zeros(ncol)
    for n in ("liq_ice_exchange", "vap_liq_exchange", "vap_ice_exchange"):
        main[n][...] = 0                                   # eamxx_p3_run.cpp:27-29
    oracle.p3_main(main, dt, predict_nc, prescribed_ccn, want_diag=False)
    post = {"pmid": fm["p_mid"], "pseudo_density": fm["pseudo_density"], "pseudo_density_dry": fm["pseudo_density_dry"],
            "th_atm": main["th"], "T_atm": fm["T_mid"], "T_prev": fm["T_prev_micro_step"], "qv_prev": fm["qv_prev_micro_step"],
            "diag_eff_radius_qc": fm["eff_radius_qc"], "diag_eff_radius_qi": fm["eff_radius_qi"],
            "diag_eff_radius_qr": fm["eff_radius_qr"], "precip_liq_surf_flux": main["precip_liq_surf"],
            "precip_ice_surf_flux": main["precip_ice_surf"], "precip_liq_surf_mass": fm["precip_liq_surf_mass"],
            "precip_ice_surf_mass": fm["precip_ice_surf_mass"]}
    for n in ("qv", "qc", "nc", "qr", "nr", "qi", "qm", "ni", "bm"):
        post[n] = fm[n]
    if fluxes:
        for n in ("vapor_flux", "water_flux", "ice_flux", "heat_flux"):
            post[n] = fm[n]
    oracle.postamble(post, dt)
    checks = oc.p3_checks({n: fm[n] for n, *_ in CHECKS}, CHECKS)
    if checks["qv"]["result"] == oc.REPAIRABLE:
        oc.repair(fm["qv"], 1e-13, 0.2)
    return checks


def test_process_driver_compiles_and_links():
    _build()
    assert os.path.exists(EXE)
    assert subprocess.run([EXE], capture_output=True).returncode == 1


def run_process_exe(tmp_path, fm, ncol, nlev, dt, nsteps, predict_nc, prescribed_ccn, fluxes, flavour):
    fin, fout = tmp_path / f"in_{flavour}.bin", tmp_path / f"out_{flavour}.bin"
    with open(fin, "wb") as fh:
        for n in NAMES2D + NAMES1D:
            fh.write(fm[n].tobytes())
    env = dict(os.environ)
    if flavour == "parity":
        # the same executable against the parity build of the library (deterministic transcendentals)
        d = tmp_path / "parity_lib"
        d.mkdir(exist_ok=True)
        link = d / "libp3b200.so"
        if not link.exists():
            os.symlink(p3.LIB_PATHS["parity"], link)
        env["LD_LIBRARY_PATH"] = str(d) + os.pathsep + env.get("LD_LIBRARY_PATH", "")
    subprocess.check_call([EXE, p3.TABLE_BLOB, str(fin), str(fout), str(ncol), str(nlev), str(dt), str(nsteps),
                           str(int(predict_nc)), str(int(prescribed_ccn)), str(int(fluxes))], env=env)
    raw = np.fromfile(fout, dtype=np.float64)
    got, off = {}, 0
    for n in NAMES2D:
        got[n] = raw[off:off + ncol * nlev].reshape(ncol, nlev); off += ncol * nlev
    for n in NAMES1D:
        got[n] = raw[off:off + ncol]; off += ncol
    return got, raw[off:off + 45].reshape(15, 3)


@pytest.mark.gpu
@pytest.mark.parametrize("predict_nc,prescribed_ccn,fluxes", [(True, True, True), (True, False, False), (False, False, True)])
def test_process_run_impl_matches_the_oracle_sequence(oracle, tmp_path, predict_nc, prescribed_ccn, fluxes):
    _build()
    ncol, nlev, dt, nsteps = 48, 72, 300.0, 2
    fm = field_manager_state(ncol, nlev, ic="mixed")
    names = [n for n in NAMES2D + NAMES1D if fluxes or n not in ("vapor_flux", "water_flux", "ice_flux", "heat_flux")]

    # (1) parity build of the library against the oracle on the same transcendentals: every field of every column
    oracle.set_math_mode("shared")
    try:
        ref = {k: v.copy() for k, v in fm.items()}
        for _ in range(nsteps):
            ref_checks = oracle_run_impl(oracle, ref, dt, predict_nc, prescribed_ccn, fluxes)
    finally:
        oracle.set_math_mode("libm")
    got, got_checks = run_process_exe(tmp_path, fm, ncol, nlev, dt, nsteps, predict_nc, prescribed_ccn, fluxes, "parity")
    errs = parity.field_errors(ref, got, names=names)
    assert max(errs.values()) <= 1e-13, {k: v for k, v in errs.items() if v > 1e-13}
    for i, (name, *_rest) in enumerate(CHECKS):
        assert int(got_checks[i, 0]) == ref_checks[name]["result"], name

    # (2) product build (libdevice) against the oracle on libm: the multi-step tolerance class on the columns the two
    # math libraries do not drive apart (a threshold flipped by a last bit changes ni by 1e-4)
    ref = {k: v.copy() for k, v in fm.items()}
    for _ in range(nsteps):
        ref_checks = oracle_run_impl(oracle, ref, dt, predict_nc, prescribed_ccn, fluxes)
    got, got_checks = run_process_exe(tmp_path, fm, ncol, nlev, dt, nsteps, predict_nc, prescribed_ccn, fluxes, "product")
    bad = np.zeros(ncol, dtype=bool)
    for n in names:
        den = np.max(np.abs(ref[n])) or 1.0
        bad |= (np.abs(ref[n] - got[n]).reshape(ncol, -1).max(axis=1) / den) > 1e-9
    print(f"product build: {bad.sum()} of {ncol} columns beyond 1e-9")
    assert bad.mean() <= 0.5
    for n in names:      # and no field drifts in the domain mean
        den = np.abs(ref[n]).mean()
        if den > 0:
            assert abs(ref[n].mean() - got[n].mean()) / den <= 1e-3, n
    for i, (name, *_rest) in enumerate(CHECKS):
        assert int(got_checks[i, 0]) == ref_checks[name]["result"], name
