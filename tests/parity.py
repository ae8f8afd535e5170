"""Shared parity machinery (tests, smoke, bench sanity): run the oracle and the CUDA path on the same
inputs and compare field by field with the reference's own metric, the max-normalised error
max|a-b| / max|a| (share/util/scream_utils.hpp:305-341 as used by p3_run_and_cmp), plus bit-exact
comparison of the per-column integer diagnostics."""
import numpy as np

from scream_b200.fields import OUTPUTS, PROGNOSTIC

# Stated tolerances (BASELINE.json north_star): 1e-12 in double, 1e-5 in float, on the max-normalised
# error of each output field after one p3_main call on identical inputs.
TOL = {8: 1e-12, 4: 1e-5}

HISTORY = ("liq_ice_exchange", "vap_liq_exchange", "vap_ice_exchange")


def clone(fields):
    return {k: v.copy() for k, v in fields.items()}


def zero_outputs(fields):
    """What P3Microphysics::run_impl does before each call (eamxx_p3_run.cpp:27-29) plus the
    pre-zeroing of p3_main_f (p3_functions_f90.cpp:1941-1945)."""
    for n in OUTPUTS:
        if n not in PROGNOSTIC:
            fields[n][...] = 0


def field_errors(ref, got, names=OUTPUTS):
    errs = {}
    for n in names:
        x = np.asarray(ref[n], dtype=np.float64)
        y = np.asarray(got[n], dtype=np.float64)
        assert x.shape == y.shape, n
        bad = ~(np.isfinite(x) & np.isfinite(y))
        if bad.any() and not np.array_equal(np.isnan(x), np.isnan(y)):
            errs[n] = np.inf
            continue
        den = np.max(np.abs(x)) if x.size else 0.0
        num = np.max(np.abs(x - y)) if x.size else 0.0
        errs[n] = num / den if den > 0 else num
    return errs


def column_mismatch(ref, got, tol, names=OUTPUTS):
    """Boolean per column: any field differs by more than tol (normalised by the field's max)."""
    ncol = ref["qc"].shape[0]
    bad = np.zeros(ncol, dtype=bool)
    for n in names:
        x = np.asarray(ref[n], dtype=np.float64)
        y = np.asarray(got[n], dtype=np.float64)
        den = np.max(np.abs(x))
        if den == 0:
            den = 1.0
        d = np.abs(x - y) / den
        bad |= (d.reshape(ncol, -1).max(axis=1) > tol)
    return bad


ULP = {8: 2.0 ** -52, 4: 2.0 ** -23}


def stable_columns(oracle, fields, dt, predict_nc, prescribed_ccn, ref=None, d_ref=None, tol=None):
    """Conditioning probe.  P3 is a cascade of thresholds (qsmall gates, int(Co_max+1) sub-step counts,
    table-cell truncation) and the reference concedes that std:: and CUDA math differ in the last bit
    (physics_share.cpp:17): for some columns that last bit decides the outcome.  This re-runs the ORACLE
    with the result of every transcendental call scaled by (1 +- 1 ulp), one function family at a time and
    all together, and returns the boolean mask of columns whose outputs stay within `tol` (default: a tenth
    of the parity tolerance) and whose integer diagnostics do not move.  Only those columns can be expected
    to agree between ANY two correct implementations that use different math libraries."""
    rb = fields["qc"].dtype.itemsize
    tol = TOL[rb] / 10 if tol is None else tol
    if ref is None:
        ref = clone(fields)
        zero_outputs(ref)
        _, d_ref = oracle.p3_main(ref, dt, predict_nc, prescribed_ccn)
    ncol = fields["qc"].shape[0]
    stable = np.ones(ncol, dtype=bool)
    n = len(oracle.MATH_FUNCS)
    probes = [np.full(n, 1 + ULP[rb]), np.full(n, 1 - ULP[rb])]
    for i in range(n):
        for sgn in (+1, -1):
            p = np.ones(n)
            p[i] = 1 + sgn * ULP[rb]
            probes.append(p)
    try:
        for p in probes:
            oracle.set_math_scale(p)
            alt = clone(fields)
            zero_outputs(alt)
            _, d_alt = oracle.p3_main(alt, dt, predict_nc, prescribed_ccn)
            stable &= ~column_mismatch(ref, alt, tol)
            stable &= ~(d_ref != d_alt).any(axis=1)
    finally:
        oracle.set_math_scale(None)
    return stable


def run_pair(oracle, handle, fields, dt, predict_nc, prescribed_ccn, nsteps=1, host_path=False):
    """Step oracle and GPU side by side from the same start; returns (ref, got, diag_ref, diag_got)."""
    from scream_b200 import lib as p3
    ref = clone(fields)
    if host_path:
        got = clone(fields)
    else:
        got = p3.to_device(fields, f"cuda:{handle.device}")
    d_ref = d_got = None
    for _ in range(nsteps):
        zero_outputs(ref)
        _, d_ref = oracle.p3_main(ref, dt, predict_nc, prescribed_ccn)
        if host_path:
            zero_outputs(got)
            _, d_got = handle.main_host(got, dt, predict_nc, prescribed_ccn)
        else:
            for n in OUTPUTS:
                if n not in PROGNOSTIC:
                    got[n].zero_()
            handle.main(got, dt, predict_nc, prescribed_ccn)
            d_got = handle.get_diag(ref["qc"].shape[0])
    if not host_path:
        got = p3.to_host(got)
    return ref, got, d_ref, d_got
