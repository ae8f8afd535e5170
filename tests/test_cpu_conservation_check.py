"""numpy restatement of MassAndEnergyColumnConservationCheck (oracle/p3_oracle_checks.py) against closed forms.

The reference ships no test for this class; the restatement is pinned on cases whose answer follows from the formulas
of mass_and_energy_column_conservation_check.cpp:283-363 by hand."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import p3_oracle_checks as oc  # noqa: E402


def state(ncol=4, nlev=8):
    f = {n: np.zeros((ncol, nlev)) for n in ("qv", "qc", "qi", "qr", "T_mid", "pseudo_density")}
    f["horiz_winds"] = np.zeros((ncol, 2, nlev))
    for n in ("ps", "phis", "vapor_flux", "water_flux", "ice_flux", "heat_flux"):
        f[n] = np.zeros(ncol)
    return f


def test_mass_and_energy_closed_forms():
    f = state()
    f["pseudo_density"][...] = oc.GRAVIT            # dp/g == 1 on every level
    f["qv"][...] = 1e-3; f["qc"][...] = 2e-3; f["qi"][...] = 3e-3; f["qr"][...] = 4e-3
    f["T_mid"][...] = 250.0
    f["horiz_winds"][:, 0, :] = 3.0; f["horiz_winds"][:, 1, :] = 4.0
    f["ps"][...] = 1.0e5; f["phis"][...] = 10.0
    m = oc.column_mass(f)
    assert np.allclose(m, 8 * 1e-2, rtol=1e-15)
    e = oc.column_energy(f)
    per_level = 250.0 * oc.CPAIR + 0.5 * 25.0 + (oc.LATVAP + oc.LATICE) * 1e-3 + oc.LATICE * (2e-3 + 4e-3)
    assert np.allclose(e, 8 * per_level + 10.0 * 1.0e5 / oc.GRAVIT, rtol=1e-15)


def test_unchanged_state_with_zero_fluxes_passes_and_flux_accounting():
    f = state()
    rng = np.random.default_rng(3)
    f["pseudo_density"][...] = rng.uniform(500, 1500, f["qv"].shape)
    for n in ("qv", "qc", "qi", "qr"):
        f[n][...] = rng.uniform(0, 1e-3, f["qv"].shape)
    f["T_mid"][...] = rng.uniform(200, 300, f["qv"].shape)
    f["ps"][...] = 1e5; f["phis"][...] = 50.0
    m0, e0 = oc.column_mass(f), oc.column_energy(f)
    r = oc.column_conservation_check(f, m0, e0, 300.0, 1e-14, 1e-14)
    assert r["pass_"] and r["mass_rel_err"] == 0.0 and r["energy_rel_err"] == 0.0
    assert r["mass_col"] == 0 and r["energy_col"] == 0        # 0 > identity: the first column holds the maximum
    # rain that left through the surface: water_flux [m/s] * RHO_H2O * dt of mass, the same mass removed from column 2
    dt, wf = 300.0, 1.0e-6
    g = {k: v.copy() for k, v in f.items()}
    g["qr"][2, -1] -= wf * oc.RHO_H2O * dt * oc.GRAVIT / g["pseudo_density"][2, -1]
    g["water_flux"][2] = wf
    r = oc.column_conservation_check(g, m0, e0, dt, 1e-12, 1.0)
    assert r["mass_rel_err"] < 1e-13
    # without the flux the check fails and points at column 2
    g["water_flux"][2] = 0.0
    r = oc.column_conservation_check(g, m0, e0, dt, 1e-12, 1.0)
    assert not r["pass_"] and r["mass_col"] == 2
    assert np.isclose(r["mass_rel_err"], wf * oc.RHO_H2O * dt / m0[2], rtol=1e-9)


def test_nan_never_wins_and_ties_report_the_first_column():
    f = state()
    f["pseudo_density"][...] = oc.GRAVIT
    f["qv"][...] = 1e-3
    f["T_mid"][...] = 250.0
    m0, e0 = oc.column_mass(f), oc.column_energy(f)
    g = {k: v.copy() for k, v in f.items()}
    g["qv"][1, 0] = np.nan
    g["qv"][3, 0] += 1e-4
    g["qv"][2, 0] += 1e-4
    r = oc.column_conservation_check(g, m0, e0, 1.0, 1e-12, 1e-12)
    assert r["mass_col"] == 2 and not r["pass_"]
