"""CPU checks of the product's column-driver SOURCE (p3_column.cuh / p3_physics.cuh) against the oracle.

The same headers the sm_100a kernel instantiates are compiled for the host with a 1-lane warp
(tests/hostsim).  These tests prove the driver logic (fusion of init/part1/part2, deferred history
stores, register-window sedimentation, lazy density factors, host-precomputed gamma values) without
a GPU; the GPU parity tests proper are in test_gpu_parity.py.
"""
import numpy as np
import pytest

from scream_b200 import ic_cases
from tests import hostsim_lib, oracle_lib, parity

FLAGS = [(False, False), (False, True), (True, False), (True, True)]


def _pair(oracle, f, dt, predict_nc, prescribed_ccn, nsteps=1):
    ref, got = parity.clone(f), parity.clone(f)
    d_ref = d_got = None
    for _ in range(nsteps):
        parity.zero_outputs(ref)
        parity.zero_outputs(got)
        _, d_ref = oracle.p3_main(ref, dt, predict_nc, prescribed_ccn)
        d_got = hostsim_lib.p3_main(got, dt, predict_nc, prescribed_ccn)
    return ref, got, d_ref, d_got


@pytest.mark.parametrize("predict_nc,prescribed_ccn", FLAGS)
@pytest.mark.parametrize("dtype,nsteps,tol", [(np.float64, 1, 1e-14), (np.float32, 1, 1e-6),
                                              (np.float64, 6, 1e-9)])
def test_mixed_ic(oracle, dtype, nsteps, tol, predict_nc, prescribed_ccn):
    """p3_run_and_cmp protocol (p3_run_and_cmp.cpp:218-352): mixed(3,72), dt=300, 1 and 6 steps.
    One step: rounding-level agreement.  Six steps: last-bit differences (the driver reuses shared
    sub-expressions) pass through the threshold cascade, hence the looser multi-step bound.  Single
    precision is only compared after one step: over several steps a last-bit difference moves a
    sub-step count (the falling front advances one level per sub-step) and the runs decorrelate."""
    f = ic_cases.make_mixed(3, 72, dtype)
    ref, got, d_ref, d_got = _pair(oracle, f, 300.0, predict_nc, prescribed_ccn, nsteps=nsteps)
    errs = parity.field_errors(ref, got)
    assert max(errs.values()) <= tol, {k: v for k, v in errs.items() if v > tol}
    assert np.array_equal(d_ref, d_got)


@pytest.mark.parametrize("nlev,dt", [(72, 1800.0), (128, 300.0), (33, 300.0), (200, 100.0)])
def test_random_columns(oracle, nlev, dt):
    """G2 random cloud layers, incl. a ragged level count and nlev > 128 (8 slots per lane)."""
    f = ic_cases.make_random(768, nlev)
    ref, got, d_ref, d_got = _pair(oracle, f, dt, True, True)
    bad = parity.column_mismatch(ref, got, 1e-13)
    assert not bad.any(), f"{bad.sum()} columns differ"
    assert np.array_equal(d_ref, d_got)


@pytest.mark.parametrize("nlev,dt", [(72, 1800.0), (128, 300.0)])
def test_product_math_on_well_conditioned_columns(oracle, nlev, dt):
    """The arithmetic the GPU build uses (x*sqrt(x) for x^1.5 etc., see Phys::pow1p5) differs from the
    oracle's libm calls by <= 1 ulp per call: compared on the columns the conditioning probe calls stable,
    to the stated tolerance 1e-12, integer diagnostics bit-exact."""
    f = ic_cases.make_random(1024, nlev)
    ref, got = parity.clone(f), parity.clone(f)
    parity.zero_outputs(ref); parity.zero_outputs(got)
    _, d_ref = oracle.p3_main(ref, dt, True, True)
    d_got = hostsim_lib.p3_main(got, dt, True, True, flavour="product")
    stable = parity.stable_columns(oracle, f, dt, True, True, ref=ref, d_ref=d_ref, tol=parity.TOL[8] / 10)
    bad = parity.column_mismatch(ref, got, parity.TOL[8])
    diag_bad = (d_ref != d_got).any(axis=1)
    assert stable.mean() > 0.5
    assert not (bad & stable).any() and not (diag_bad & stable).any()


def test_clear_sky_early_exit(oracle):
    """Columns with no hydrometeors and no nucleation leave after part1 (p3_main_impl.hpp:237):
    defaults for the effective radii, history arrays untouched, zero precipitation."""
    f = ic_cases.make_mixed(4, 72)
    for n in ("qc", "nc", "qr", "nr", "qi", "qm", "ni", "bm"):
        f[n][...] = 0
    f["qv"][...] = 1e-7
    f["th"][...] = 300.0 * f["inv_exner"]      # warm: no nucleation possible
    ref, got, d_ref, d_got = _pair(oracle, f, 300.0, True, True)
    assert (d_ref[:, :2] == 0).all()
    assert np.array_equal(d_ref, d_got)
    for n in parity.OUTPUTS:
        assert np.array_equal(ref[n], got[n]), n
    assert (got["diag_eff_radius_qc"] == 10e-6).all() and (got["precip_liq_surf"] == 0).all()


def test_single_level_and_single_column(oracle):
    f = ic_cases.make_random(1, 1)
    ref, got, d_ref, d_got = _pair(oracle, f, 300.0, True, True)
    for n in parity.OUTPUTS:
        np.testing.assert_allclose(got[n], ref[n], rtol=1e-13, atol=0, err_msg=n)
    assert np.array_equal(d_ref, d_got)


@pytest.mark.parametrize("nlev,dt,dtype", [(72, 1800.0, np.float64), (128, 300.0, np.float64), (128, 300.0, np.float32)])
def test_parity_flavour_is_bit_identical_to_the_shared_math_oracle(nlev, dt, dtype):
    """-DP3B200_PARITY (reference operation order + p3_math.hpp) vs the oracle switched to the same header: every
    output of every random column bit for bit, integer diagnostics included — the CPU twin of
    tests/test_gpu_parity_all_columns.py."""
    from scream_b200.fields import OUTPUTS
    oracle_lib.set_math_mode("shared")
    try:
        f = ic_cases.make_random(1024, nlev, dtype=dtype)
        parity.zero_outputs(f)
        ref, got = parity.clone(f), parity.clone(f)
        _, d_ref = oracle_lib.p3_main(ref, dt, True, True)
        d_got = hostsim_lib.p3_main(got, dt, True, True, flavour="parity")
    finally:
        oracle_lib.set_math_mode("libm")
    for n in OUTPUTS:
        assert np.array_equal(ref[n], got[n], equal_nan=True), n
    assert np.array_equal(d_ref[:, :12], d_got[:, :12])


@pytest.mark.parametrize("flavour", ["exact", "product", "parity"])
@pytest.mark.parametrize("nlev,dt", [(72, 1800.0), (128, 300.0)])
def test_rolled_ice_window_is_bit_identical_to_the_unrolled_one(monkeypatch, flavour, nlev, dt):
    """Sediment::run_rolled (ice on tall windows: state in an array, slots as a rolled loop — what
    p3_sediment<ice, 4> runs on the device) against Sediment::run (slots unrolled in registers): every output and the
    sub-step counts bit for bit, on columns with ice windows of all heights.  Here a slot is one level (warp of size
    1), so every level boundary exercises the carry between slots and the slot the front enters."""
    from scream_b200.fields import OUTPUTS
    f = ic_cases.make_random(512, nlev)
    parity.zero_outputs(f)
    a, b = parity.clone(f), parity.clone(f)
    monkeypatch.setenv("P3B200_HOSTSIM_ICE", "unrolled")
    d_a = hostsim_lib.p3_main(a, dt, True, True, flavour=flavour)
    monkeypatch.setenv("P3B200_HOSTSIM_ICE", "rolled")
    d_b = hostsim_lib.p3_main(b, dt, True, True, flavour=flavour)
    assert (d_a[:, 9] >= 0).sum() > 100          # D_QI_KTOP: most columns sediment ice
    assert d_a[:, 11].max() > 1                   # D_QI_NSUB: several sub-steps somewhere
    for n in OUTPUTS:
        assert np.array_equal(a[n], b[n], equal_nan=True), n
    assert np.array_equal(d_a, d_b)
