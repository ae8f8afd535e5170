"""The reference's own closed-form PROPERTY tests, restated as assertions against the oracle.

The reference stores no expected outputs for p3_main; what pins its functions are
the property tests of p3/tests/*_unit_tests.cpp.  Each test below cites the reference test it restates and
drives the oracle through oracle/p3_oracle_unit.cpp (the analogue of the reference's *_f bridges).
"""
import ctypes

import numpy as np
import pytest

D = ctypes.c_double
P = ctypes.POINTER(ctypes.c_double)
I = ctypes.c_int


def _ptr(a):
    return a.ctypes.data_as(P)


@pytest.fixture(scope="module")
def unit(oracle):
    L = oracle.lib()
    L.p3o_unit_upwind2.argtypes = [I, I, I, I, D, P, P, P, P, P, P]
    L.p3o_unit_upwind2.restype = None
    L.p3o_unit_generalized_sedimentation1.argtypes = [I, I, I, I, I, D, P, P, P, P, P, P, P, P]
    L.p3o_unit_generalized_sedimentation1.restype = I
    for f in (L.p3o_unit_find_top, L.p3o_unit_find_bottom):
        f.argtypes = [P, D, I, I, I, ctypes.POINTER(I)]
        f.restype = I
    for f in (L.p3o_unit_cloud_water_conservation, L.p3o_unit_rain_water_conservation, L.p3o_unit_ice_water_conservation):
        f.argtypes = [D, D, P]
        f.restype = None
    L.p3o_unit_evaporate_rain.argtypes = [P, P, P]
    L.p3o_unit_evaporate_rain.restype = None
    L.p3o_unit_rain_evap_tscale_weight.argtypes = [D]
    L.p3o_unit_rain_evap_tscale_weight.restype = D
    L.p3o_unit_qv_sat_dry.argtypes = [D, D, I]
    L.p3o_unit_qv_sat_dry.restype = D
    L.p3o_unit_apply_table_ice.argtypes = [P, I, D, D, D, D]
    L.p3o_unit_apply_table_ice.restype = D
    return L


# order of Procs<Real>::Tend (oracle/p3_oracle_procs.hpp)
TEND = ["qc2qr_accret", "qr2qv_evap", "qc2qr_autoconv", "nc_accret", "nc_selfcollect", "nc2nr_autoconv",
        "nr_selfcollect", "nr_evap", "ncautr", "qi2qv_sublim", "nr_ice_shed", "qc2qi_hetero_freeze", "qr2qi_collect",
        "qc2qr_ice_shed", "qi2qr_melt", "qc2qi_collect", "qr2qi_immers_freeze", "qv2qi_nucleat", "ni2nr_melt",
        "nc_collect", "ncshdc", "nc2ni_immers_freeze", "nr_collect", "ni_selfcollect", "ni_nucleat", "qv2qi_vapdep",
        "qc2qi_berg", "nr2ni_immers_freeze", "ni_sublim", "qc_growth_rate"]
TI = {n: i for i, n in enumerate(TEND)}


def _tend(**kw):
    t = np.zeros(30)
    for k, v in kw.items():
        t[TI[k]] = v
    return t


@pytest.mark.parametrize("nk", [17, 32, 77, 128])
@pytest.mark.parametrize("kdir", [-1, 1])
def test_upwind_conserves_mass_and_is_monotone(unit, nk, kdir):
    """p3_upwind_unit_tests.cpp:39-200 (TestUpwind::run_phys): 2*nk first-order upwind steps at CFL <= 1 on
    a stretched grid.  Mass leaves only through the outflow face (mass + accumulated boundary flux is
    conserved), the true mixing ratio r1/r0 never exceeds its previous maximum, and nothing goes negative."""
    eps = np.finfo(np.float64).eps
    max_speed, min_dz = 4.2, 0.33
    dt = min_dz / max_speed
    k = np.arange(nk, dtype=np.float64)
    rho = 1 + k / nk
    inv_rho = 1 / rho
    inv_dz = 1 / (min_dz + k * k / (nk * nk))
    V = np.stack([0.5 * (1 + k / nk) * max_speed] * 2)
    r = np.stack([np.ones(nk), np.where((k >= 2) & (k < nk - 2), k / nk, 0.0)])
    flux = np.zeros((2, nk))
    k_bot, k_top = (0, nk - 1) if kdir == 1 else (nk - 1, 0)
    r_max0 = (r[1] / r[0]).max()
    for _ in range(2 * nk):
        mass0 = (rho * r[1] / inv_dz).sum()
        unit.p3o_unit_upwind2(nk, k_bot, k_top, kdir, dt, _ptr(rho), _ptr(inv_rho), _ptr(inv_dz), _ptr(flux), _ptr(V), _ptr(r))
        mass1 = (rho * r[1] / inv_dz).sum()
        # mass change == what left through the bottom face
        assert abs((mass0 - mass1) - flux[1, k_bot] * dt) <= 1e3 * eps * max(mass0, 1e-300)
        assert (r >= 0).all()
        ok = eps * r[0] >= eps                       # skip cells where r0 has been advected to noise
        r_max1 = (r[1][ok] / r[0][ok]).max() if ok.any() else 0.0
        assert r_max1 <= r_max0 * (1 + 1e1 * eps) + 1e1 * eps
        r_max0 = max(r_max1, 0.0) if ok.any() else r_max0


def test_generalized_sedimentation_substep_and_front(unit):
    """p3_upwind_impl.hpp:97-129: dt_sub = dt_left / int(Co_max + 1), the front k_qxbot advances one level per
    sub-step until it reaches kbot, surface flux accumulates only once it has."""
    nk = 20
    rho = np.ones(nk); inv_rho = np.ones(nk); inv_dz = np.full(nk, 0.01)
    r = np.zeros(nk); r[5] = 1.0
    V = np.full(nk, 3.0); flux = np.zeros(nk)
    dt_left = D(100.0); prt = D(0.0)
    kb = unit.p3o_unit_generalized_sedimentation1(nk, 5, 5, nk - 1, -1, 2.7, ctypes.byref(dt_left), ctypes.byref(prt),
                                                  _ptr(rho), _ptr(inv_rho), _ptr(inv_dz), _ptr(flux), _ptr(V), _ptr(r))
    assert kb == 6 and prt.value == 0.0
    assert dt_left.value == 100.0 - 100.0 / 3           # int(2.7 + 1) = 3
    assert r[6] > 0 and abs(r.sum() - 1.0) < 1e-15      # mass moved one level down, none lost


def test_find_top_bottom(unit):
    """p3_find_unit_tests.cpp:27-101: first/last level with v >= small, both directions, and the empty case."""
    nk = 37
    small = 0.5
    for kdir in (-1, 1):
        kbot, ktop = (nk - 1, 0) if kdir == -1 else (0, nk - 1)
        pres = I(0)
        for lo, hi in [(3, 20), (0, nk - 1), (17, 17)]:
            v = np.zeros(nk); v[lo:hi + 1] = 1.0
            kt = unit.p3o_unit_find_top(_ptr(v), small, kbot, ktop, kdir, ctypes.byref(pres))
            assert pres.value == 1 and kt == (lo if kdir == -1 else hi)
            kb = unit.p3o_unit_find_bottom(_ptr(v), small, kbot, kt, kdir, ctypes.byref(pres))
            assert pres.value == 1 and kb == (hi if kdir == -1 else lo)
        v = np.full(nk, 0.1)
        unit.p3o_unit_find_top(_ptr(v), small, kbot, ktop, kdir, ctypes.byref(pres))
        assert pres.value == 0


def test_cloud_water_conservation(unit):
    """p3_unit_tests.cpp:24-70 (TestP3Conservation::cloud_water_conservation_tests_device): sinks are only
    rescaled when they exceed the available cloud water, and then they sum to exactly qc/dt."""
    dt = 1.1
    t = _tend(qc2qr_autoconv=1e-4, qc2qr_accret=0.0, qc2qi_collect=0.0, qc2qi_hetero_freeze=0.0, qc2qr_ice_shed=0.0,
              qc2qi_berg=0.0, qi2qv_sublim=1.0, qv2qi_vapdep=1.0)
    before = t.copy()
    unit.p3o_unit_cloud_water_conservation(1e-5, dt, _ptr(t))
    ratio = 1e-5 / (1e-4 * dt)
    assert abs(t[TI["qc2qr_autoconv"]] * dt - 1e-5) < 1e-19
    assert t[TI["qi2qv_sublim"]] == before[TI["qi2qv_sublim"]] * (1 - ratio)   # l.43-47: split by (1-ratio)
    t = _tend(qc2qr_autoconv=1e-9)
    unit.p3o_unit_cloud_water_conservation(1e-4, dt, _ptr(t))                  # sinks << sources: untouched
    assert t[TI["qc2qr_autoconv"]] == 1e-9
    t = _tend(qc2qr_autoconv=0.0)
    unit.p3o_unit_cloud_water_conservation(0.0, dt, _ptr(t))                   # nothing to scale
    assert (t == 0).all()


def test_rain_and_ice_water_conservation(unit):
    """p3_unit_tests.cpp:72-175: same property for rain (sources include the conversions into rain) and ice."""
    dt = 1.1
    t = _tend(qr2qv_evap=2e-5, qr2qi_collect=1e-5, qr2qi_immers_freeze=1e-5, qc2qr_autoconv=1e-6)
    unit.p3o_unit_rain_water_conservation(1e-5, dt, _ptr(t))
    sinks = (t[TI["qr2qv_evap"]] + t[TI["qr2qi_collect"]] + t[TI["qr2qi_immers_freeze"]]) * dt
    assert abs(sinks - (1e-5 + 1e-6 * dt)) < 1e-19
    t0 = _tend(qr2qv_evap=1e-9)
    t = t0.copy()
    unit.p3o_unit_rain_water_conservation(1e-4, dt, _ptr(t))
    assert np.array_equal(t, t0)
    t = _tend(qi2qv_sublim=3e-5, qi2qr_melt=1e-5, qv2qi_vapdep=2e-6)
    unit.p3o_unit_ice_water_conservation(1e-5, dt, _ptr(t))
    sinks = (t[TI["qi2qv_sublim"]] + t[TI["qi2qr_melt"]]) * dt
    assert abs(sinks - (1e-5 + 2e-6 * dt)) < 1e-19


def test_rain_evaporation_limits(unit):
    """p3_evaporate_rain_unit_tests.cpp:24-140 (TestEvapSublPrecip::run_property)."""
    w = unit.p3o_unit_rain_evap_tscale_weight
    assert 0 <= w(1e-8) <= 1 and 1 - w(1e-8) < 1e-8
    assert 0 <= w(1e15) <= 1 and w(1e15) < 1e-8
    base = dict(qr_incld=1e-4, qc_incld=1e-4, nr_incld=1e7, qi_incld=1e-4, cld_frac_l=0.1, cld_frac_r=0.9, qv=5e-4,
                qv_prev=1e-3, qv_sat_l=1e-3, qv_sat_i=6e-4, ab=1.2, abi=1.2, epsr=1 / 20, epsi_tot=1 / 60, t=287.0,
                t_prev=285.0, dqsdt=1e-3, dt=60.0)

    def evap(**kw):
        a = dict(base); a.update(kw)
        arr = np.array(list(a.values()))
        q, n = D(0), D(0)
        unit.p3o_unit_evaporate_rain(_ptr(arr), ctypes.byref(q), ctypes.byref(n))
        return q.value, n.value, a

    q, n, _ = evap(qr_incld=1e-14 / 2)                       # below QSMALL: no evaporation
    assert abs(q) < 1e-8 and abs(n) < 1e-8
    q, n, a = evap(qr_incld=5e-13)                           # tiny rain evaporates within the step
    assert abs(q - 5e-13 / a["dt"] * (a["cld_frac_r"] - a["cld_frac_l"]) / a["cld_frac_r"]) < 1e-8
    assert n <= a["nr_incld"] / a["dt"]
    q, n, _ = evap(cld_frac_r=0.1, cld_frac_l=0.9)           # no rainy area outside cloud
    assert q == 0 and n == 0
    q, n, _ = evap(qv=2e-3)                                  # supersaturated: no evaporation
    assert q == 0 and n == 0
    q, n, a = evap()                                         # generic case: bounded by the rain that is there
    assert 0 <= q <= a["qr_incld"] / a["dt"]
    q, n, a = evap(t=260.0, t_prev=259.0)                    # cold branch stays bounded too
    assert 0 <= q <= a["qr_incld"] / a["dt"]


def test_saturation_is_monotone_and_matches_at_melting(unit):
    """physics_saturation_impl.hpp:41-83: Murphy-Koop liquid and ice vapour pressures coincide near the triple
    point (to 0.1 %), ice < liquid below it, both increase with T; qv_sat scales as 1/p."""
    qs = unit.p3o_unit_qv_sat_dry
    T = np.linspace(190.0, 310.0, 200)
    liq = np.array([qs(t, 1e5, 0) for t in T])
    ice = np.array([qs(t, 1e5, 1) for t in T])
    assert (np.diff(liq) > 0).all() and (np.diff(ice) > 0).all()
    assert (ice[T < 273.0] < liq[T < 273.0]).all()
    assert abs(qs(273.16, 1e5, 0) / qs(273.16 - 1e-9, 1e5, 1) - 1) < 1e-3
    assert abs(qs(280.0, 5e4, 0) / qs(280.0, 1e5, 0) - 2) < 1e-12


def test_ice_table_interpolation_reproduces_linear_tables(unit):
    """p3_ice_tables_unit_tests.cpp:65-97 (scaffolded there, disabled with #if 0 at :299): a table that is
    linear in each index is reproduced exactly by the trilinear interpolation of apply_table_ice, and the
    interpolated value equals the linear function of the continuous coordinates lookup_ice forms."""
    jj, ii, i = np.meshgrid(np.arange(5), np.arange(4), np.arange(50), indexing="ij")
    table = np.zeros((5, 4, 50, 12))
    for idx in range(12):
        table[..., idx] = 3.0 * jj + 7.0 * ii + 0.5 * i + idx
    rng = np.random.default_rng(7)
    c = 4.135985029041767
    for _ in range(200):
        qi = 10 ** rng.uniform(-8, -3); ni = 10 ** rng.uniform(2, 7)
        qm = qi * rng.uniform(0, 1); rhop = rng.uniform(50, 900)
        dum1 = min(max((np.log10(qi / ni) + 18) * c - 10, 1.0), 50.0)
        dum4 = min(max(qm / qi * 3 + 1, 1.0), 4.0)
        dum5 = (rhop - 50) * 0.005 + 1 if rhop <= 650 else (rhop - 650) * 0.004 + 4
        dum5 = min(max(dum5, 1.0), 5.0)
        for idx in (0, 5, 11):
            got = unit.p3o_unit_apply_table_ice(_ptr(table), idx, qi, ni, qm, rhop)
            want = 3.0 * (dum5 - 1) + 7.0 * (dum4 - 1) + 0.5 * (dum1 - 1) + idx
            assert abs(got - want) < 1e-11, (qi, ni, qm, rhop, idx, got, want)
