"""The 15 property checks of the P3 process as one fused reduction (p3_checks.cuh; C ABI p3b200_postamble_checked,
p3b200_property_checks, p3b200_repair_qv) against the numpy oracle (oracle/p3_oracle_checks.py).

min/max are selections, not arithmetic: on the same array they are BIT-exact, so are the locations (lowest flat index)
and the Pass/Repairable/Fail classification.  The fused postamble carries a violation mask only: when all 15 checks
pass it returns no extreme values (NaN), otherwise it continues with the full reduction.  For the fused postamble the checked values are the ones the kernel has
just written, so the oracle is applied to the kernel's own outputs (which test_gpu_prepost.py ties to the oracle's
postamble) and, separately, the classification is compared with the oracle-only sequence."""
import numpy as np
import pytest

from oracle import p3_oracle_checks as oc
from scream_b200 import prepost_cases
from scream_b200.fields import CHECK_COLUMN_FIELDS, CHECK_FIELDS, CHECK_POSTAMBLE_SOURCE, CHECKS

pytestmark = pytest.mark.gpu

KEYS = ("lb", "ub", "min_val", "max_val", "result", "pass_lower", "pass_upper")


def _to_dev(torch, f):
    return {k: torch.from_numpy(v).cuda() for k, v in f.items()}


def _compare(got, ref, values=True):
    for name in CHECK_FIELDS:
        g, r = got[name], ref[name]
        for k in KEYS:
            if k in ("min_val", "max_val") and not values:
                assert np.isnan(g[k]), (name, k, g[k])     # all 15 passed: the fused call carries no extrema
                continue
            assert g[k] == r[k], (name, k, g[k], r[k])
        if r["result"] == oc.PASS:
            assert g["min_loc"] == g["max_loc"] == -1, name      # locations are only looked up for a failing check
        else:
            assert (g["min_loc"], g["max_loc"]) == (r["min_loc"], r["max_loc"]), name


def _check_inputs(fields):
    """{check name: array} from postamble-named arrays"""
    return {n: fields[CHECK_POSTAMBLE_SOURCE.get(n, n)] for n in CHECK_FIELDS}


def _violate(f, dtype):
    """Plant the violations the reference's checks exist for, before the postamble (values scale through it)."""
    f["qc"][5, 17] = 0.35            # > 0.1 after the dry->wet rescale as well: Fail, upper
    f["qc"][400, 100] = 0.35         # a second, equal maximum further on: the lowest index is reported
    for i, k in ((5, 17), (400, 100)):
        f["pseudo_density_dry"][i, k] = f["pseudo_density"][i, k]     # same rescale factor (1) for the two
    f["ni"][9, 0] = -3.0             # Fail, lower
    f["qv"][11, 127] = 1e-15         # below 1e-13: Repairable (qv is registered with can_repair)
    f["bm"][0, 0] = np.nan           # NaN never wins a comparison
    f["diag_eff_radius_qi"][77, 3] = 6e-3          # 6000 micron > 5000
    f["precip_ice_surf_mass"][123] = -1.0          # accumulated mass below 0 even after this step's flux
    f["precip_ice_surf_flux"][123] = 0.0


@pytest.mark.parametrize("rb", [8, 4])
@pytest.mark.parametrize("violate", [False, True])
def test_postamble_checked_matches_oracle(oracle, p3lib, rb, violate):
    torch = pytest.importorskip("torch")
    dtype = np.float64 if rb == 8 else np.float32
    f = prepost_cases.make_postamble(613, 128, dtype, with_fluxes=True)
    if violate:
        _violate(f, dtype)
    h = p3lib.P3Handle(0, rb)
    plain = _to_dev(torch, f)
    h.postamble(plain, 300.0)
    dev = _to_dev(torch, f)
    got = h.postamble_checked(dev, 300.0)
    out = {k: v.cpu().numpy() for k, v in dev.items()}
    for k in out:                                    # the checked variant writes exactly what the plain one writes
        assert np.array_equal(out[k], plain[k].cpu().numpy(), equal_nan=True), k
    _compare(got, oc.p3_checks(_check_inputs(out), CHECKS), values=violate)
    # the classification of the oracle-only sequence (oracle postamble -> oracle checks)
    ref = {k: v.copy() for k, v in f.items()}
    oracle.postamble(ref, 300.0)
    ref_checks = oc.p3_checks(_check_inputs(ref), CHECKS)
    for n in CHECK_FIELDS:
        assert got[n]["result"] == ref_checks[n]["result"], n
        assert (got[n]["min_loc"], got[n]["max_loc"]) == ((ref_checks[n]["min_loc"], ref_checks[n]["max_loc"])
                                                          if ref_checks[n]["result"] != oc.PASS else (-1, -1)), n
    if violate:
        res = {n: got[n]["result"] for n in CHECK_FIELDS}
        assert res["qc"] == res["ni"] == res["eff_radius_qi"] == res["precip_ice_surf_mass"] == oc.FAIL
        assert res["qv"] == oc.REPAIRABLE and res["bm"] == oc.PASS and res["T_mid"] == oc.PASS
        assert got["qc"]["max_loc"] == 5 * 128 + 17 and got["ni"]["min_loc"] == 9 * 128
        assert got["precip_ice_surf_mass"]["min_loc"] == 123
        h.repair_qv(dev["qv"])
        qv = out["qv"].copy()
        assert np.array_equal(dev["qv"].cpu().numpy(), oc.repair(qv, 1e-13, 0.2))
        again = h.property_checks(_check_inputs(dev))
        # fp32: float(1e-13) lies just below the double bound 1e-13 the check compares against, so the clipped field
        # is still "repairable" — the reference's repair_impl (ST lb = m_lb) and check_impl behave the same way
        assert again["qv"]["result"] == oc.check(qv, 1e-13, 0.2, True)["result"] == (oc.PASS if rb == 8 else oc.REPAIRABLE)
        assert again["qc"]["result"] == oc.FAIL
    else:
        assert all(got[n]["result"] == oc.PASS for n in CHECK_FIELDS), got
    h.close()


@pytest.mark.parametrize("rb,ncol,nlev,padded", [(8, 1, 72, False), (8, 1031, 128, True), (4, 257, 72, False)])
def test_standalone_checks_bit_exact(p3lib, rb, ncol, nlev, padded):
    torch = pytest.importorskip("torch")
    dtype = np.float64 if rb == 8 else np.float32
    rng = np.random.default_rng(ncol)
    stride = nlev + 8 if padded else nlev
    host, dev = {}, {}
    for i, (name, lb, ub, _) in enumerate(CHECKS):
        shape = (ncol,) if name in CHECK_COLUMN_FIELDS else (ncol, nlev)
        hi = min(ub, 1e12)
        a = rng.uniform(lb, hi, shape).astype(dtype)
        a = np.clip(a, dtype(lb), dtype(hi)) if name != "qv" else np.clip(a, dtype(1e-12), dtype(0.19))
        if i % 3 == 0:
            a.reshape(-1)[rng.integers(a.size)] = dtype(-7.5)       # below every lower bound
        if i % 4 == 1:
            a.reshape(-1)[rng.integers(a.size)] = dtype(3e12)       # above every finite upper bound
        host[name] = a
        if a.ndim == 2:
            buf = torch.zeros((ncol, stride), dtype=torch.float64 if rb == 8 else torch.float32, device="cuda")
            buf[:, nlev:] = 1e30                                    # padding must not be read
            buf[:, :nlev] = torch.from_numpy(a).cuda()
            dev[name] = buf[:, :nlev]
        else:
            dev[name] = torch.from_numpy(a).cuda()
    h = p3lib.P3Handle(0, rb)
    _compare(h.property_checks(dev), oc.p3_checks(host, CHECKS))
    h.close()


def test_full_size_checks_pass_after_p3(p3lib):
    """ne120pg2-sized grid: preamble-free sequence p3_main -> property checks on the prognostic outputs; the extrema
    of the tiled grid equal the extrema of its 96 distinct columns."""
    torch = pytest.importorskip("torch")
    from scream_b200 import ic_cases
    from scream_b200.fields import FIELDS
    from tests import parity
    ncol, nlev, tile_n = 345600, 128, 96
    tile = ic_cases.make_mixed_tiled(tile_n, nlev, tile=3, dtype=np.float64)
    parity.zero_outputs(tile)
    idx = torch.arange(ncol, device="cuda") % tile_n
    fields = {n: torch.from_numpy(tile[n]).cuda().index_select(0, idx).contiguous() for n in FIELDS}
    h = p3lib.P3Handle(0, 8)
    h.main(fields, 300.0, True, True)
    chk = {"T_mid": fields["th"] / fields["inv_exner"], "qv": fields["qv"], "qc": fields["qc"], "qi": fields["qi"],
           "qr": fields["qr"], "qm": fields["qm"], "nc": fields["nc"], "nr": fields["nr"], "ni": fields["ni"],
           "bm": fields["bm"], "precip_liq_surf_mass": fields["precip_liq_surf"] * 300.0 * 1000.0,
           "precip_ice_surf_mass": fields["precip_ice_surf"] * 300.0 * 1000.0,
           "eff_radius_qc": fields["diag_eff_radius_qc"] * 1e6, "eff_radius_qi": fields["diag_eff_radius_qi"] * 1e6,
           "eff_radius_qr": fields["diag_eff_radius_qr"] * 1e6}
    chk = {k: v.contiguous() for k, v in chk.items()}
    got = h.property_checks(chk)
    ref = oc.p3_checks({k: v[:tile_n].cpu().numpy() for k, v in chk.items()}, CHECKS)
    for n in CHECK_FIELDS:
        for k in ("min_val", "max_val", "result"):
            assert got[n][k] == ref[n][k], (n, k, got[n][k], ref[n][k])
    h.close()
