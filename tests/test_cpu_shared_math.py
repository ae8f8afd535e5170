"""Accuracy of the deterministic transcendentals shared by the kernels and the oracle (scream_b200/csrc/p3_math.hpp).

They replace libm / libdevice so that the CPU oracle and the GPU agree bit for bit (the reference does the same
between its Fortran and C++ builds, physics/share/physics_share.cpp:17-150).  Here their error against mpmath
(50 digits) is measured in ulps over the argument ranges P3 feeds them, plus the IEEE special cases.
"""
import math

import mpmath as mp
import numpy as np
import pytest

from tests import oracle_lib

mp.mp.dps = 50


def ulp_err(got, exact):
    """|got - exact| in units of the last place of `exact` (binary64)."""
    errs = []
    for g, e in zip(got, exact):
        if e == 0:
            errs.append(0.0 if g == 0 else float("inf"))
            continue
        ef = float(e)
        u = math.ulp(ef) if math.isfinite(ef) and ef != 0 else 5e-324
        errs.append(float(abs(mp.mpf(float(g)) - e) / u))
    return np.array(errs)


RNG = np.random.default_rng(20241105)
N = 4000

CASES = [
    # name, sampler of x (and y), mpmath function, max ulp bound
    ("exp", lambda: (RNG.uniform(-700, 700, N), None), lambda x, y: mp.exp(x), 1.0),
    ("exp", lambda: (RNG.uniform(-2, 2, N), None), lambda x, y: mp.exp(x), 1.0),
    ("log", lambda: (np.exp(RNG.uniform(-700, 700, N)), None), lambda x, y: mp.log(x), 1.0),
    ("log", lambda: (RNG.uniform(0.5, 2.0, N), None), lambda x, y: mp.log(x), 1.0),
    ("log10", lambda: (np.exp(RNG.uniform(-100, 100, N)), None), lambda x, y: mp.log10(x), 1.0),
    ("pow", lambda: (np.exp(RNG.uniform(-50, 50, N)), RNG.uniform(-12, 12, N)), lambda x, y: mp.power(x, y), 1.5),
    ("pow", lambda: (RNG.uniform(150, 320, N), np.full(N, 1.81)), lambda x, y: mp.power(x, y), 1.5),     # T^1.81
    ("pow", lambda: (RNG.uniform(0.2, 30, N), np.full(N, 0.54)), lambda x, y: mp.power(x, y), 1.5),      # rhofac
    ("pow", lambda: (np.exp(RNG.uniform(-20, -4, N)), np.full(N, 2.47)), lambda x, y: mp.power(x, y), 1.5),  # autoconv
    ("pow", lambda: (np.full(N, 10.0), RNG.uniform(-40, 40, N)), lambda x, y: mp.power(x, y), 1.5),      # 10^x
    ("cbrt", lambda: (np.exp(RNG.uniform(-700, 700, N)) * RNG.choice([-1, 1], N), None), lambda x, y: mp.cbrt(abs(x)) * (1 if x > 0 else -1), 1.0),
    ("expm1", lambda: (RNG.uniform(-40, 40, N), None), lambda x, y: mp.expm1(x), 2.0),
    ("expm1", lambda: (RNG.uniform(-1, 1, N) * np.exp(RNG.uniform(-30, 0, N)), None), lambda x, y: mp.expm1(x), 2.0),
    ("tanh", lambda: (RNG.uniform(-25, 25, N), None), lambda x, y: mp.tanh(x), 2.5),
    ("tanh", lambda: (RNG.uniform(-1, 1, N), None), lambda x, y: mp.tanh(x), 2.5),
    ("tgamma", lambda: (RNG.uniform(1, 25, N), None), lambda x, y: mp.gamma(x), 5.0),       # P3: mu_c + 1..7, mu_r + 1..7
    ("tgamma", lambda: (RNG.uniform(0.01, 171, N), None), lambda x, y: mp.gamma(x), 6.0),
]


@pytest.mark.parametrize("idx", range(len(CASES)))
def test_ulp_error_against_mpmath(idx):
    name, sampler, fn, bound = CASES[idx]
    x, y = sampler()
    got = oracle_lib.math_eval(name, x, y)
    exact = [fn(mp.mpf(float(a)), mp.mpf(float(b)) if y is not None else None)
             for a, b in zip(x, y if y is not None else x)]
    e = ulp_err(got, exact)
    print(f"{name}: max {e.max():.3f} ulp, mean {e.mean():.3f} ulp over {len(x)} points")
    assert e.max() <= bound, (name, e.max(), x[e.argmax()])


def test_special_values():
    ev = oracle_lib.math_eval
    inf, nan = float("inf"), float("nan")
    assert ev("exp", [800.0, -800.0, inf, -inf, 0.0]).tolist() == [inf, 0.0, inf, 0.0, 1.0]
    assert np.isnan(ev("exp", [nan])[0])
    r = ev("log", [0.0, -1.0, inf, 1.0])
    assert r[0] == -inf and np.isnan(r[1]) and r[2] == inf and r[3] == 0.0
    r = ev("log10", [0.0, 100.0, 1e-5, 1.0])
    assert r[0] == -inf and r[1] == 2.0 and r[2] == -5.0 and r[3] == 0.0
    r = ev("pow", [2.0, -8.0, 0.0, 0.0, 5.0, 1.0, -8.0, 3.0, 3.0], [0.5, 3.0, -1.0, 2.0, 0.0, nan, 0.5, 2.0, 3.0])
    assert abs(r[0] - math.sqrt(2.0)) <= math.ulp(1.0) and r[1] == -512.0 and r[2] == inf and r[3] == 0.0 and r[4] == 1.0 and r[5] == 1.0
    assert np.isnan(r[6]) and r[7] == 9.0 and r[8] == 27.0
    r = ev("cbrt", [27.0, -27.0, 0.0, inf, 1e-320, 8e-300])
    assert r[0] == 3.0 and r[1] == -3.0 and r[2] == 0.0 and r[3] == inf
    assert abs(r[4] - 1e-320 ** (1 / 3)) < 1e-120 and abs(r[5] / 2e-100 - 1) < 1e-15
    r = ev("tgamma", [1.0, 2.0, 5.0, 0.5, 171.0, 172.0, -1.0, -0.5, 10.0])
    assert r[0] == 1.0 and r[1] == 1.0 and r[2] == 24.0 and abs(r[3] - math.sqrt(math.pi)) < 4e-16
    assert math.isfinite(r[4]) and r[5] == inf and np.isnan(r[6]) and abs(r[7] + 2 * math.sqrt(math.pi)) < 2e-15
    assert r[8] == 362880.0
    r = ev("expm1", [0.0, 1e-20, -800.0, 800.0])
    assert r.tolist() == [0.0, 1e-20, -1.0, inf]
    r = ev("tanh", [0.0, 30.0, -30.0, 1e-10])
    assert r.tolist() == [0.0, 1.0, -1.0, 1e-10]


def test_float_face_is_the_rounded_double():
    x = RNG.uniform(0.1, 20, 1000).astype(np.float32).astype(np.float64)
    for name in ("exp", "log", "log10", "cbrt", "tgamma", "tanh", "expm1"):
        d = oracle_lib.math_eval(name, x)
        f = oracle_lib.math_eval(name, x, real_bytes=4)
        assert np.array_equal(d.astype(np.float32).astype(np.float64), f), name
