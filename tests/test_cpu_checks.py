"""The numpy oracle of the P3 property checks (oracle/p3_oracle_checks.py) against the reference's own known-answer
cases for FieldWithinIntervalCheck / FieldLowerBoundCheck (share/tests/property_checks.cpp, field of extents (2,3,12)),
and the table of P3's 15 registered checks against eamxx_p3_process_interface.cpp:226-243."""
import numpy as np

from oracle import p3_oracle_checks as oc
from scream_b200.fields import CHECKS, CHECK_FIELDS

SHAPE = (2, 3, 12)


def _flat(*idx):
    return int(np.ravel_multi_index(idx, SHAPE))


def test_interval_check_known_answers():
    # property_checks.cpp:137-176: interval [0,1], can_repair, default repairable bounds
    rng = np.random.default_rng(0)
    f = rng.uniform(0.01, 0.99, SHAPE)
    assert oc.check(f, 0, 1, True)["result"] == oc.PASS
    f[...] = 0.5
    f[1, 2, 3] = 2.0
    f[0, 1, 2] = 0.0
    r = oc.check(f, 0, 1, True)
    assert r["result"] == oc.REPAIRABLE
    assert (r["min_val"], r["min_loc"]) == (0.0, _flat(0, 1, 2))
    assert (r["max_val"], r["max_loc"]) == (2.0, _flat(1, 2, 3))
    oc.repair(f, 0, 1)
    assert oc.check(f, 0, 1, True)["result"] == oc.PASS


def test_interval_check_repairable_bounds_known_answers():
    # property_checks.cpp:179-250: interval [0,1], repairable bounds [-1,2]
    rng = np.random.default_rng(1)
    f = rng.uniform(0.01, 0.99, SHAPE)
    flat = f.reshape(-1)
    flat[0], flat[-1] = -0.5, 1.5
    assert oc.check(f, 0, 1, True, -1, 2)["result"] == oc.REPAIRABLE
    flat[0], flat[-1] = -2, 3
    r = oc.check(f, 0, 1, True, -1, 2)
    assert r["result"] == oc.FAIL
    assert (r["min_val"], r["min_loc"]) == (-2.0, _flat(0, 0, 0))
    assert (r["max_val"], r["max_loc"]) == (3.0, _flat(1, 2, 11))
    oc.repair(f, 0, 1)
    assert oc.check(f, 0, 1, True, -1, 2)["result"] == oc.PASS
    flat[1] = -2                      # lower bound only: the failure location is the minimum's
    r = oc.check(f, 0, 1, True, -1, 2)
    assert r["result"] == oc.FAIL and r["fail_loc"] == _flat(0, 0, 1) and (r["pass_lower"], r["pass_upper"]) == (0, 1)
    oc.repair(f, 0, 1)
    flat[-1] = 4                      # upper bound only
    r = oc.check(f, 0, 1, True, -1, 2)
    assert r["result"] == oc.FAIL and r["fail_loc"] == _flat(1, 2, 11) and (r["pass_lower"], r["pass_upper"]) == (1, 0)


def test_lower_bound_check_known_answers():
    # property_checks.cpp:253-290: FieldLowerBoundCheck(-1, can_repair): values near the largest double pass
    n = int(np.prod(SHAPE))
    f = (np.finfo(np.float64).max - np.arange(n, dtype=np.float64)).reshape(SHAPE)
    assert oc.check(f, -1.0, oc.S_MAX, True)["result"] == oc.PASS
    f = -np.arange(n, dtype=np.float64).reshape(SHAPE)
    r = oc.check(f, -1.0, oc.S_MAX, True)
    assert r["result"] == oc.REPAIRABLE and r["min_loc"] == n - 1
    oc.repair(f, -1.0, oc.S_MAX)
    assert oc.check(f, -1.0, oc.S_MAX, True)["result"] == oc.PASS


def test_nan_is_skipped_and_identities():
    f = np.full((4, 5), np.nan)
    r = oc.check(f, 0.0, 1.0)
    big = np.finfo(np.float64).max
    # the identities satisfy lb <= min and max <= ub: a field of NaNs passes an interval check in the reference
    # (catching NaNs is FieldNaNCheck's job, not this one's)
    assert (r["min_val"], r["max_val"]) == (big, -big) and r["result"] == oc.PASS
    f[2, 3] = 0.25
    r = oc.check(f, 0.0, 1.0)
    assert r["result"] == oc.PASS and r["min_val"] == r["max_val"] == 0.25
    g = np.array([np.nan, 2.0, -1.0], dtype=np.float32)
    assert np.array_equal(oc.repair(g, 0, 1), np.array([np.nan, 1.0, 0.0], dtype=np.float32), equal_nan=True)


def test_registered_checks_table():
    # eamxx_p3_process_interface.cpp:226-243, in order
    assert CHECK_FIELDS == ("T_mid", "qv", "qc", "qi", "qr", "qm", "nc", "nr", "ni", "bm", "precip_liq_surf_mass",
                            "precip_ice_surf_mass", "eff_radius_qc", "eff_radius_qi", "eff_radius_qr")
    table = {c[0]: c[1:] for c in CHECKS}
    assert table["T_mid"] == (100.0, 500.0, False) and table["qv"] == (1e-13, 0.2, True)
    for n in ("qc", "qi", "qr", "qm"):
        assert table[n] == (0.0, 0.1, False)
    assert table["nc"] == (0.0, 1e11, False) and table["nr"] == table["ni"] == (0.0, 1e10, False)
    assert table["bm"] == (0.0, 1.0, False)
    assert table["precip_liq_surf_mass"] == table["precip_ice_surf_mass"] == (0.0, oc.S_MAX, False)
    assert table["eff_radius_qc"] == (0.0, 100.0, False)
    assert table["eff_radius_qi"] == table["eff_radius_qr"] == (0.0, 5000.0, False)
