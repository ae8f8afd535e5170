"""Synthetic P3 input columns (numpy, host side; used by tests and bench.py).

G0/G1: `make_mixed` restates the reference's only shipped initial condition,
  ic::Factory::mixed (components/eamxx/src/physics/p3/p3_ic_cases.cpp:11-120), column by column.
  For throughput grids the 3 distinct column flavours (i % 3 temperature perturbation) are tiled:
  the reference formula `plo = max(i, ...)` at k=0 makes dz negative for i > ~500, so big grids use
  column (i % tile) of a small `tile`-column case instead of i itself.
G2: `make_random` — seeded per-column random cloud layers (value ranges of
  run_bfb_p3_main, p3_main_unit_tests.cpp:397-421, on a physical T/p profile).
"""
import numpy as np

from .fields import FIELDS, INTERFACE_FIELDS, SURFACE_1D

# physics_constants.hpp:27-98
P0, RD, CP = 100000.0, 287.042, 1004.64


def empty_fields(ncol, nlev, dtype=np.float64):
    d = {}
    for n in FIELDS:
        if n in SURFACE_1D:
            d[n] = np.zeros((ncol,), dtype)
        elif n in INTERFACE_FIELDS:
            d[n] = np.zeros((ncol, nlev + 1), dtype)
        else:
            d[n] = np.zeros((ncol, nlev), dtype)
    return d


def make_mixed(ncol, nlev, dtype=np.float64):
    """ic::Factory::mixed for columns i = 0..ncol-1 (p3_ic_cases.cpp:11-120), computed in double."""
    nk = nlev
    d = empty_fields(ncol, nlev, np.float64)
    k = np.arange(nk, dtype=np.float64)
    i = np.arange(ncol, dtype=np.float64)[:, None]
    k15 = np.arange(15, dtype=np.float64)
    k20 = np.arange(20, dtype=np.float64)
    # integer divisions i/ncol and k/nk are both 0 in the reference (l.27)
    d["nccn"][:] = (100.0 + 0 - 0) * 1e6
    d["qc"][:, nk - 20:nk - 5] = 1e-4 * (1 - k15 / 14)
    d["nc"][:] = 1e6
    d["qr"][:, nk - 20:nk] = 1e-5 * (1 - k20 / 19)
    d["nr"][:] = 1e6
    d["qi"][:, nk - 20:nk - 5] = 1e-4
    d["ni"][:] = 1e6
    d["qm"][:, nk - 20:nk - 5] = 1e-4 * (1 - k15 / 14)
    d["bm"][:, nk - 20:nk - 5] = 1e-2
    tmp = -5e-4 + 1e-3 / float(nk) * k
    d["qv"][:] = np.where(tmp > 0, tmp, 0.0)
    d["qv"][:, nk - 20:nk - 5] = 5e-3
    pres = 100 + 1e5 / float(nk) * k
    d["pres"][:] = pres
    d["dpres"][:] = 1e5 / float(nk)
    d["inv_exner"][:] = np.power(1e5 / pres, 287.15 / 1005.0)
    for n in ("cld_frac_i", "cld_frac_l", "cld_frac_r", "inv_qc_relvar"):
        d[n][:] = 1.0
    T = np.broadcast_to(150 + 150 / float(nk) * k, (ncol, nk)).copy()
    pert = ((np.arange(ncol) % 3) - 0.5)[:, None] / float(nk) * k
    pert[0, :] = 0.0                       # `if (i > 0)` (l.72)
    T = T + pert
    d["th"][:] = T * np.power(P0 / pres, RD / CP)
    d["qi"][:, nk - 1] = 1e-9
    d["qv"][:, nk - 1] = 5e-2
    d["qr"][:, nk - 1] = 1e-6
    d["qc"][:, nk - 1] = 1e-6
    d["qi"][:, nk - 25] = 5e-8
    d["qc"][:, 35] = 1e-7
    d["qv"][:, 35] = 1e-6
    d["qv"][:, 33] = 1e-4
    d["qv_prev"][:] = d["qv"]
    d["t_prev"][:] = T
    g = 9.8
    plo = np.empty((ncol, nk))
    phi = np.empty((ncol, nk))
    plo[:, 1:] = 0.5 * (pres[:-1] + pres[1:])
    plo[:, 0] = np.maximum(i[:, 0], pres[0] - 0.5 * (pres[1] - pres[0]) / (1 - 0))
    phi[:, :-1] = 0.5 * (pres[:-1] + pres[1:])
    phi[:, -1] = pres[-1] + 0.5 * (pres[-1] - pres[-2]) / (1 - 0)
    d["dz"][:] = RD * T / (g * pres) * (phi - plo)
    if dtype != np.float64:
        d = {n: v.astype(dtype) for n, v in d.items()}
    return d


def drop_cold_cloud(fields, t_min=235.0):
    """Single-precision variant of `mixed`: cloud water at T < t_min removed.  With 128 levels the reference IC puts a
    cloudy level at T = 191 K, where Bigg immersion freezing evaluates exp(0.65*(273.15 - T)) = 1.5e23 times gamma
    factors: the REFERENCE ITSELF overflows binary32 there (its own single-precision build returns inf/NaN in qv and th
    at that level, oracle/_ref sp build) — the case is outside what P3 can do in single precision."""
    T = fields["th"].astype(np.float64) / fields["inv_exner"].astype(np.float64)
    cold = T < t_min
    for n in ("qc", "nc"):
        fields[n][cold] = 0
    return fields


def make_mixed_tiled(ncol, nlev, tile=3, dtype=np.float64, single_precision_safe=False):
    """G1: column i := mixed(tile, nlev)[i % tile]."""
    base = make_mixed(tile, nlev, dtype)
    if single_precision_safe:
        base = drop_cold_cloud(base)
    idx = np.arange(ncol) % tile
    return {n: np.ascontiguousarray(v[idx]) for n, v in base.items()}


def make_random(ncol, nlev, seed=20241104, dtype=np.float64, p_cloud=0.5):
    """G2: per-column random activity on a physical background profile."""
    rng = np.random.Generator(np.random.MT19937(seed))
    nk = nlev
    d = empty_fields(ncol, nlev, np.float64)
    k = np.arange(nk, dtype=np.float64)
    pres = 100 + 1e5 / float(nk) * k
    d["pres"][:] = pres
    d["dpres"][:] = 1e5 / float(nk)
    d["inv_exner"][:] = np.power(P0 / pres, RD / CP)
    T = 190.0 + 110.0 * (k / float(nk - 1)) + rng.uniform(-3.0, 3.0, size=(ncol, 1))
    d["th"][:] = T * d["inv_exner"]
    d["t_prev"][:] = T + rng.uniform(-0.2, 0.2, size=(ncol, nk))
    g = 9.80616
    d["dz"][:] = RD * T / (g * pres) * d["dpres"]
    # Murphy-Koop liquid saturation (physics_saturation_impl.hpp:62-71) for a qv background
    logt = np.log(T)
    es = np.exp(54.842763 - 6763.22 / T - 4.210 * logt + 0.000367 * T +
                np.tanh(0.0415 * (T - 218.8)) * (53.878 - 1331.22 / T - 9.44523 * logt + 0.014025 * T))
    qs = 0.621971966 * es / np.maximum(pres, 1e-3)
    d["qv"][:] = qs * rng.uniform(0.3, 1.1, size=(ncol, nk))
    d["qv_prev"][:] = d["qv"] * rng.uniform(0.98, 1.02, size=(ncol, nk))
    for n in ("cld_frac_i", "cld_frac_l", "cld_frac_r"):
        d[n][:] = 1e-4
    d["inv_qc_relvar"][:] = 1.0
    d["nccn"][:] = 100e6
    d["nc_nuceat_tend"][:] = rng.uniform(0.0, 1e3, size=(ncol, nk))
    d["ni_activated"][:] = rng.uniform(0.0, 1e4, size=(ncol, nk))
    has_cloud = rng.uniform(size=ncol) < p_cloud
    thick = rng.integers(4, 41, size=ncol)
    lo = int(0.4 * nk)
    top = rng.integers(lo, np.maximum(nk - thick, lo + 1))
    kk = np.arange(nk)[None, :]
    mask = has_cloud[:, None] & (kk >= top[:, None]) & (kk < (top + thick)[:, None])
    u = lambda a, b: rng.uniform(a, b, size=(ncol, nk))
    d["qc"][:] = np.where(mask, u(0, 1e-4), 0.0)
    d["qr"][:] = np.where(mask, u(0, 1e-5), 0.0)
    d["qi"][:] = np.where(mask, u(0, 1e-4), 0.0)
    d["qm"][:] = np.where(mask, u(0, 1.0) * d["qi"], 0.0)
    d["bm"][:] = np.where(mask, d["qm"] / u(100.0, 800.0), 0.0)
    for n in ("nc", "nr", "ni"):
        d[n][:] = 1e6
    for n in ("cld_frac_i", "cld_frac_l"):
        d[n][:] = np.where(mask, u(0.1, 1.0), 1e-4)
    # max-overlap rain fraction, as p3_preamble builds it (eamxx_p3_process_interface.hpp:112-127)
    cf = np.maximum(d["cld_frac_l"], d["cld_frac_i"])
    d["cld_frac_r"][:] = np.maximum.accumulate(cf, axis=1)
    if dtype != np.float64:
        d = {n: v.astype(dtype) for n, v in d.items()}
    return d
