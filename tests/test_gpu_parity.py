"""GPU parity tests proper: the sm_100a kernel through the C ABI vs the CPU oracle."""
import numpy as np
import pytest

from scream_b200 import ic_cases
from scream_b200.fields import OUTPUTS, PROGNOSTIC
from tests import parity

pytestmark = pytest.mark.gpu

FLAGS = [(False, False), (False, True), (True, False), (True, True)]


@pytest.fixture(scope="module")
def handle64(p3lib):
    h = p3lib.P3Handle(0, 8)
    yield h
    h.close()


@pytest.fixture(scope="module")
def handle32(p3lib):
    h = p3lib.P3Handle(0, 4)
    yield h
    h.close()


def _report(errs, tol):
    return {k: v for k, v in errs.items() if v > tol}


@pytest.mark.parametrize("predict_nc,prescribed_ccn", FLAGS)
def test_mixed_ic_one_step_double(oracle, handle64, predict_nc, prescribed_ccn):
    """G0: ic::Factory::mixed(3, 72), dt=300 — the p3_run_and_cmp case (p3_run_and_cmp.cpp:218-352)."""
    f = ic_cases.make_mixed(3, 72)
    ref, got, d_ref, d_got = parity.run_pair(oracle, handle64, f, 300.0, predict_nc, prescribed_ccn)
    errs = parity.field_errors(ref, got)
    assert not _report(errs, parity.TOL[8]), _report(errs, parity.TOL[8])
    assert np.array_equal(d_ref, d_got)
    assert handle64.error_summary() == 0


@pytest.mark.parametrize("predict_nc,prescribed_ccn", FLAGS)
def test_mixed_ic_six_steps_double(oracle, handle64, predict_nc, prescribed_ccn):
    """p3_run_and_cmp protocol: 6 consecutive steps; rounding differences may be amplified by the
    threshold cascade, so the multi-step tolerance is the reference's own baseline tolerance class
    (p3_run_and_cmp.cpp: tol = 0 only for BFB builds; otherwise 1e-5..) — we keep 1e-9."""
    f = ic_cases.make_mixed(3, 72)
    ref, got, d_ref, d_got = parity.run_pair(oracle, handle64, f, 300.0, predict_nc, prescribed_ccn, nsteps=6)
    errs = parity.field_errors(ref, got)
    assert not _report(errs, 1e-9), _report(errs, 1e-9)
    assert np.array_equal(d_ref, d_got)


def _stable_parity(oracle, handle, f, dt, rb, min_stable_frac):
    """Parity on the columns the conditioning probe (tests/parity.py: stable_columns) calls well
    conditioned: there the kernel must be within the stated tolerance and the integer diagnostics
    bit-exact.  The remaining columns are decided by the last bit of libm vs libdevice (the oracle itself
    moves by more than the tolerance when its transcendentals are nudged by 1 ulp); they are counted and
    printed, never hidden, and compared in aggregate."""
    ref, got, d_ref, d_got = parity.run_pair(oracle, handle, f, dt, True, True)
    stable = parity.stable_columns(oracle, f, dt, True, True, ref=ref, d_ref=d_ref, tol=parity.TOL[rb] / 10)
    bad = parity.column_mismatch(ref, got, parity.TOL[rb])
    diag_bad = (d_ref != d_got).any(axis=1)
    print(f"stable columns {stable.sum()} / {stable.size}; over tolerance: {bad.sum()} (of which stable "
          f"{(bad & stable).sum()}); diag mismatches: {diag_bad.sum()} (of which stable {(diag_bad & stable).sum()})")
    assert stable.mean() >= min_stable_frac
    assert not (bad & stable).any(), f"{(bad & stable).sum()} well-conditioned columns exceed the tolerance"
    assert not (diag_bad & stable).any()
    # aggregate agreement over ALL columns: domain means of the prognostic state and of surface precipitation
    # (precipitation is the small residue of the ill-conditioned columns: looser bound)
    agg = {8: 1e-3, 4: 2e-2}[rb]
    for n, tol in (("qc", agg), ("qr", agg), ("qi", agg), ("qv", agg), ("th", agg),
                   ("precip_liq_surf", 50 * agg), ("precip_ice_surf", 50 * agg)):
        x, y = float(np.mean(ref[n], dtype=np.float64)), float(np.mean(got[n], dtype=np.float64))
        assert abs(x - y) <= tol * max(abs(x), 1e-300), (n, x, y)
    assert handle.error_summary() == 0


@pytest.mark.parametrize("nlev,dt", [(72, 1800.0), (128, 300.0)])
def test_random_columns_double(oracle, handle64, nlev, dt):
    """G2 random activity, 4096 columns, tolerance 1e-12."""
    _stable_parity(oracle, handle64, ic_cases.make_random(4096, nlev), dt, 8, 0.5)


def test_random_columns_float(oracle, handle32):
    """Single precision, tolerance 1e-5 on the well-conditioned columns.  In float every column that
    holds hydrometeors is ill conditioned in the sense above (a 1-ulp nudge of the oracle's own
    transcendentals moves it by more than 1e-5: qsmall = 1e-14 sits at the edge of float resolution of
    the tendencies), so the per-column check covers the clear-sky half and the aggregate check the rest."""
    _stable_parity(oracle, handle32, ic_cases.make_random(4096, 128, dtype=np.float32), 300.0, 4, 0.3)


@pytest.mark.parametrize("predict_nc,prescribed_ccn", FLAGS)
def test_mixed_ic_float_aggregate(oracle, handle32, predict_nc, prescribed_ccn):
    """ic::Factory::mixed in float: all three columns are ill conditioned in float (see above); the
    kernel must still run clean and agree with the float oracle in the column-integrated sense."""
    f = ic_cases.make_mixed(3, 72, np.float32)
    ref, got, d_ref, d_got = parity.run_pair(oracle, handle32, f, 300.0, predict_nc, prescribed_ccn)
    assert handle32.error_summary() == 0
    assert np.array_equal(d_ref[:, :3], d_got[:, :3])
    w = f["dpres"].astype(np.float64)
    for n in ("qc", "qr", "qi", "qv"):
        x = (ref[n].astype(np.float64) * w).sum(); y = (got[n].astype(np.float64) * w).sum()
        assert abs(x - y) <= 2e-2 * abs(x), (n, x, y)
    np.testing.assert_allclose(got["th"], ref["th"], rtol=1e-4)


def test_host_path_matches_device_path(oracle, handle64):
    """p3b200_main_host (chunked pipeline) must equal p3b200_main bit for bit."""
    f = ic_cases.make_random(3000, 72)
    _, got_d, _, dd = parity.run_pair(oracle, handle64, f, 300.0, True, False)
    _, got_h, _, dh = parity.run_pair(oracle, handle64, f, 300.0, True, False, host_path=True)
    for n in OUTPUTS:
        assert np.array_equal(got_d[n], got_h[n]), n
    assert np.array_equal(dd, dh)


def test_host_path_many_chunks(handle64, p3lib, monkeypatch):
    """The host pipeline with 1 MB chunks (1024 columns, ragged last chunk, every slot reused several times) equals
    the device path bit for bit."""
    monkeypatch.setenv("P3B200_HOST_CHUNK_MB", "1")
    f = ic_cases.make_random(9973, 72)
    parity.zero_outputs(f)
    dev = p3lib.to_device(parity.clone(f))
    handle64.main(dev, 300.0, True, True)
    dd = handle64.get_diag(9973)
    want = p3lib.to_host(dev)
    got = parity.clone(f)
    _, dh = handle64.main_host(got, 300.0, True, True)
    for n in OUTPUTS:
        assert np.array_equal(want[n], got[n]), n
    assert np.array_equal(dd, dh)


def test_strided_tracer_views(oracle, handle64, p3lib):
    """Tracer fields are sub-views of a bundled (ncol, ntracer, nlev) array: column stride > nlev
    (share/field/field_impl.hpp:880-909)."""
    import torch
    f = ic_cases.make_mixed(7, 72)
    ref = parity.clone(f)
    parity.zero_outputs(ref)
    oracle.p3_main(ref, 300.0, True, True)
    dev = p3lib.to_device(f)
    tracers = ["qc", "nc", "qr", "nr", "qi", "qm", "ni", "bm", "qv"]
    bundle = torch.zeros((7, len(tracers), 72), dtype=torch.float64, device="cuda:0")
    for j, n in enumerate(tracers):
        bundle[:, j, :] = dev[n]
        dev[n] = bundle[:, j, :]
    handle64.main(dev, 300.0, True, True)
    got = p3lib.to_host(dev)
    errs = parity.field_errors(ref, got)
    assert not _report(errs, parity.TOL[8]), _report(errs, parity.TOL[8])


def test_runtime_knobs_generic_rain_table_path(oracle, handle64, p3lib):
    """Non-default P3Runtime: a non-integral constant_mu_rain makes the mu_r axis of the rain tables matter
    (bilinear path of p3_table3_impl.hpp:70-83 instead of the shared-memory slice), plus changed
    autoconversion/accretion exponents and ice sedimentation factor."""
    over = {"constant_mu_rain": 1.5, "autoconversion_qc_exponent": 2.3, "accretion_qr_exponent": 1.1,
            "ice_sedimentation_factor": 0.9, "max_total_ni": 500.0e3}
    f = ic_cases.make_mixed(3, 72)
    ref = parity.clone(f)
    parity.zero_outputs(ref)
    _, d_ref = oracle.p3_main(ref, 300.0, True, True, runtime=over)
    dev = p3lib.to_device(f)
    for n in OUTPUTS:
        if n not in PROGNOSTIC:
            dev[n].zero_()
    handle64.main(dev, 300.0, True, True, runtime=p3lib.make_runtime(over))
    got = p3lib.to_host(dev)
    errs = parity.field_errors(ref, got)
    assert not _report(errs, parity.TOL[8]), _report(errs, parity.TOL[8])
    assert np.array_equal(d_ref, handle64.get_diag(3))


@pytest.mark.parametrize("rb", [8, 4])
def test_zero_history_replaces_the_callers_zero_fill(p3lib, rb):
    """P3B200Infra::zero_history = 1 on arrays full of garbage == the caller's zero-fill (eamxx_p3_run.cpp:27-29)
    followed by the reference-shaped call, bit for bit, on every output."""
    import torch
    h = p3lib.P3Handle(0, rb)
    f = ic_cases.make_random(1024, 128, seed=7)
    if rb == 4:
        f = {k: v.astype(np.float32) for k, v in f.items()}
    a = parity.clone(f)
    parity.zero_outputs(a)
    da = p3lib.to_device(a)
    h.main(da, 300.0, True, True)
    ra = p3lib.to_host(da)
    b = parity.clone(f)
    parity.zero_outputs(b)
    for n in parity.HISTORY:
        b[n][...] = 12345.0
    db = p3lib.to_device(b)
    h.main(db, 300.0, True, True, zero_history=True)
    rbk = p3lib.to_host(db)
    for n in OUTPUTS:
        assert np.array_equal(ra[n], rbk[n], equal_nan=True), n
    h.close()
