"""ORACLE — TEST INFRASTRUCTURE ONLY.  Never imported by the product (scream_b200/).

numpy restatement of the property checks P3Microphysics registers in initialize_impl
(components/eamxx/src/physics/p3/eamxx_p3_process_interface.cpp:226-243) as the atmosphere driver evaluates them:

  FieldWithinIntervalCheck::check_impl   share/property_checks/field_within_interval_check.cpp:80-226
      min/max(+loc) over the whole field with `v < min` / `v > max` updates (NaN never wins either comparison),
      starting from Kokkos' MinMaxLoc identities (largest finite value of the field type / its negative);
      Pass when lb <= min and max <= ub; Fail when min < lb_repairable or max > ub_repairable; else Repairable.
  FieldWithinIntervalCheck constructor   :12-68      the repairable bounds are [lb, ub] unless can_repair, in which
                                                     case they default to -+ the largest double (hpp:33-35)
  FieldLowerBoundCheck                   share/property_checks/field_lower_bound_check.hpp:13-26 (ub = DBL_MAX)
  FieldWithinIntervalCheck::repair_impl  :338-372    v = min(ub, v); v = max(lb, v) with ekat's `a < b ? a : b`

Pinned against the reference's own known-answer cases for these classes (share/tests/property_checks.cpp:137-176 and
:179-230: a 0.5-filled field with one 2.0 and one 0.0, the slightly / completely out-of-bounds pair with repairable
bounds [-1, 2]), restated in tests/test_cpu_checks.py.  Where several entries hold the extreme value the reference's
location depends on Kokkos' join order; this restatement (and the CUDA path) reports the lowest flat index.
"""
import numpy as np

PASS, REPAIRABLE, FAIL = 0, 1, 2
S_MAX = np.finfo(np.float64).max


def minmaxloc(v):
    """(min, min_loc, max, max_loc) of the flattened field with the reference's comparisons."""
    flat = np.asarray(v).reshape(-1)
    big = float(np.finfo(flat.dtype).max)
    ok = ~np.isnan(flat)
    mn, mx, mn_loc, mx_loc = big, -big, -1, -1
    if ok.any():
        lo, hi = float(flat[ok].min()), float(flat[ok].max())
        if lo < mn:
            mn, mn_loc = lo, int(np.flatnonzero(flat == flat.dtype.type(lo))[0])
        if hi > mx:
            mx, mx_loc = hi, int(np.flatnonzero(flat == flat.dtype.type(hi))[0])
    return mn, mn_loc, mx, mx_loc


def check(v, lb, ub, can_repair=False, lb_repairable=-S_MAX, ub_repairable=S_MAX):
    mn, mn_loc, mx, mx_loc = minmaxloc(v)
    if can_repair:   # constructor :46-68
        assert lb_repairable <= lb and ub_repairable >= ub
    lb_rep, ub_rep = (lb_repairable, ub_repairable) if can_repair else (lb, ub)
    pass_lower = pass_upper = 1
    if mn >= lb and mx <= ub:
        result = PASS
    elif mn < lb_rep or mx > ub_rep:
        result = FAIL
        pass_lower, pass_upper = int(not mn < lb_rep), int(not mx > ub_rep)
    else:
        result = REPAIRABLE
        pass_lower, pass_upper = int(not mn < lb), int(not mx > ub)
    # :244-251 — the location reported as THE failure: the minimum's if the lower test failed, else the maximum's
    fail_loc = -1 if result == PASS else (mn_loc if not pass_lower else mx_loc)
    return {"fail_loc": fail_loc, "lb": lb, "ub": ub, "min_val": mn, "max_val": mx, "min_loc": mn_loc, "max_loc": mx_loc, "result": result,
            "pass_lower": pass_lower, "pass_upper": pass_upper}


def repair(v, lb, ub):
    """In-place clip; NaN stays NaN exactly as `ub < v ? ub : v` leaves it."""
    t = v.dtype.type
    lb, ub = t(lb), t(ub)
    v[...] = np.where(ub < v, ub, v)
    v[...] = np.where(lb > v, lb, v)
    return v


def p3_checks(fields, checks):
    """fields: {name: array}; checks: scream_b200.fields.CHECKS -> {name: check dict}"""
    return {name: check(fields[name], lb, ub, rep) for name, lb, ub, rep in checks}


# ---------------------------------------------------------------------------------------------------------------
# MassAndEnergyColumnConservationCheck  share/property_checks/mass_and_energy_column_conservation_check.cpp
#   compute_total_mass_on_column :283-301, compute_mass_boundary_flux_on_column :304-312,
#   compute_total_energy_on_column :314-347, compute_energy_boundary_flux_on_column :349-363, check() :110-215.
# Constants: physics/share/physics_constants.hpp (gravit 9.80616, Cpair 1004.64, LatVap 2501000, LatIce 333700,
# RHO_H2O 1000).  The within-column sums run level by level in index order (the reference serialises them under
# ekatBFB); everything is evaluated in the precision of the fields, as the reference does with Real.
GRAVIT, CPAIR, LATVAP, LATICE, RHO_H2O = 9.80616, 1004.64, 2501000.0, 333700.0, 1000.0


def _seq_sum(terms):
    acc = np.zeros(terms.shape[0], dtype=terms.dtype)
    for k in range(terms.shape[1]):
        acc = acc + terms[:, k]
    return acc


def column_mass(f):
    t = f["qv"].dtype.type
    terms = (f["qv"] + f["qc"] + f["qi"] + f["qr"]) * f["pseudo_density"] / t(GRAVIT)
    return _seq_sum(terms)


def column_energy(f):
    t = f["qv"].dtype.type
    u, v = f["horiz_winds"][:, 0, :], f["horiz_winds"][:, 1, :]
    u2, v2 = u * u, v * v
    terms = (f["T_mid"] * t(CPAIR) + t(0.5) * (u2 + v2) + (t(LATVAP) + t(LATICE)) * f["qv"] +
             t(LATICE) * (f["qc"] + f["qr"])) * f["pseudo_density"] / t(GRAVIT)
    return _seq_sum(terms) + f["phis"] * f["ps"] / t(GRAVIT)


def column_conservation_check(f, prev_mass, prev_energy, dt, mass_tol, energy_tol):
    """check(): returns dict(mass_rel_err, mass_col, energy_rel_err, energy_col, pass_) + the per-column errors."""
    t = f["qv"].dtype.type
    dt = t(dt)
    tm, te = column_mass(f), column_energy(f)
    tm_exp = prev_mass + (f["vapor_flux"] - f["water_flux"] * t(RHO_H2O)) * dt
    te_exp = prev_energy + (f["vapor_flux"] * (t(LATVAP) + t(LATICE)) -
                            (f["water_flux"] - f["ice_flux"]) * t(RHO_H2O) * t(LATICE) + f["heat_flux"]) * dt
    with np.errstate(divide="ignore", invalid="ignore"):
        rel_m = np.abs(tm - tm_exp) / prev_mass
        rel_e = np.abs(te - te_exp) / prev_energy

    def maxloc(v):   # Kokkos::MaxLoc identity, `v > best` updates: NaN never wins, first (lowest) column on ties
        best, loc = -float(np.finfo(v.dtype).max), -1
        for i, x in enumerate(v):
            if x > best:
                best, loc = float(x), i
        return best, loc

    mv, ml = maxloc(rel_m)
    ev, el = maxloc(rel_e)
    return {"mass_rel_err": mv, "mass_col": ml, "energy_rel_err": ev, "energy_col": el,
            "pass_": bool(mv < mass_tol and ev < energy_tol), "rel_mass": rel_m, "rel_energy": rel_e}
