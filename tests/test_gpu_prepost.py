"""p3_preamble / p3_postamble CUDA kernels (scream_b200/csrc/p3_prepost.cu, through the C ABI) against the oracle
restatement.  Only pow() in exner_function differs between libm and libdevice (<= 2 ulp), every other operation is
IEEE: tolerance 1e-14 (fp64) / 1e-6 (fp32) on the max-normalised error, arrays without pow bit-exact."""
import numpy as np
import pytest

from scream_b200 import prepost_cases
from scream_b200.fields import POSTAMBLE_OUTPUTS, PREAMBLE_OUTPUTS
from tests import parity

pytestmark = pytest.mark.gpu

EXACT_PRE = ("qv", "qc", "nc", "qr", "nr", "qi", "qm", "ni", "bm", "qv_prev", "cld_frac_l", "cld_frac_i", "cld_frac_r", "dz")
EXACT_POST = ("qv", "qc", "nc", "qr", "nr", "qi", "qm", "ni", "bm", "qv_prev", "diag_eff_radius_qc", "diag_eff_radius_qi",
              "diag_eff_radius_qr", "precip_liq_surf_mass", "precip_ice_surf_mass", "vapor_flux", "water_flux", "ice_flux",
              "heat_flux")


def _to_dev(torch, f):
    return {k: torch.from_numpy(v).cuda() for k, v in f.items()}


@pytest.mark.parametrize("rb,ncol,nlev", [(8, 517, 128), (8, 3, 72), (4, 517, 128), (8, 37, 71), (4, 37, 71)])
def test_preamble_matches_oracle(oracle, p3lib, rb, ncol, nlev):
    torch = pytest.importorskip("torch")
    dtype = np.float64 if rb == 8 else np.float32
    f = prepost_cases.make_preamble(ncol, nlev, dtype)
    ref = {k: v.copy() for k, v in f.items()}
    oracle.preamble(ref)
    h = p3lib.P3Handle(0, rb)
    dev = _to_dev(torch, f)
    h.preamble(dev)
    got = {k: v.cpu().numpy() for k, v in dev.items()}
    errs = parity.field_errors(ref, got, names=PREAMBLE_OUTPUTS)
    tol = 1e-14 if rb == 8 else 1e-6
    assert max(errs.values()) <= tol, {k: v for k, v in errs.items() if v > tol}
    for n in EXACT_PRE:
        assert np.array_equal(ref[n], got[n]), n
    for n in ("pmid", "pseudo_density", "pseudo_density_dry", "T_atm", "cld_frac_t"):
        assert np.array_equal(f[n], got[n]), n
    h.close()


@pytest.mark.parametrize("rb,with_fluxes,nlev", [(8, True, 128), (8, False, 128), (4, True, 128), (8, True, 71), (4, True, 71)])
def test_postamble_matches_oracle(oracle, p3lib, rb, with_fluxes, nlev):
    torch = pytest.importorskip("torch")
    dtype = np.float64 if rb == 8 else np.float32
    f = prepost_cases.make_postamble(613, nlev, dtype, with_fluxes=with_fluxes)
    ref = {k: v.copy() for k, v in f.items()}
    oracle.postamble(ref, 300.0)
    h = p3lib.P3Handle(0, rb)
    dev = _to_dev(torch, f)
    h.postamble(dev, 300.0)
    got = {k: v.cpu().numpy() for k, v in dev.items()}
    names = [n for n in POSTAMBLE_OUTPUTS if n in f]
    errs = parity.field_errors(ref, got, names=names)
    tol = 1e-14 if rb == 8 else 1e-6
    assert max(errs.values()) <= tol, {k: v for k, v in errs.items() if v > tol}
    for n in EXACT_POST:
        if n in f:
            assert np.array_equal(ref[n], got[n]), n
    h.close()


def test_run_impl_sequence_conserves_wet_water(oracle, p3lib):
    """preamble -> p3_main -> postamble on the GPU, the order of P3Microphysics::run_impl: runs through the three
    entry points on shared device arrays and checks the result against the same sequence on the oracle."""
    torch = pytest.importorskip("torch")
    from scream_b200 import ic_cases
    from scream_b200.fields import FIELDS, OUTPUTS, PROGNOSTIC
    ncol, nlev, dt = 96, 72, 300.0
    pre = prepost_cases.make_preamble(ncol, nlev, np.float64, seed=7)
    base = ic_cases.make_random(ncol, nlev, dtype=np.float64)
    # field-manager state: wet mixing ratios = the random case's values, thermodynamic state from the random case
    pd = base["dpres"].copy()
    pre.update(pmid=base["pres"].copy(), pseudo_density=pd, T_atm=(base["th"] / base["inv_exner"]).copy(),
               pseudo_density_dry=pd * (1 - base["qv"]))
    for n in ("qv", "qc", "nc", "qr", "nr", "qi", "qm", "ni", "bm", "qv_prev"):
        pre[n] = base[n].copy()

    def sequence(run_pre, run_main, run_post, to_np):
        f = run_pre({k: v.copy() for k, v in pre.items()})
        main = {n: base[n].copy() for n in FIELDS}
        parity.zero_outputs(main)
        fn = to_np(f)
        for n in ("qv", "qc", "nc", "qr", "nr", "qi", "qm", "ni", "bm", "qv_prev", "inv_exner", "cld_frac_l", "cld_frac_i",
                  "cld_frac_r", "dz"):
            main[n] = fn[n].copy()
        main["th"] = fn["th_atm"].copy()
        main = run_main(main)
        post = prepost_cases.make_postamble(ncol, nlev, np.float64, seed=7)
        for n in ("pmid", "pseudo_density", "pseudo_density_dry", "T_atm"):
            post[n] = pre[n].copy()
        for n in ("qv", "qc", "nc", "qr", "nr", "qi", "qm", "ni", "bm", "diag_eff_radius_qc", "diag_eff_radius_qi",
                  "diag_eff_radius_qr"):
            post[n] = main[n].copy()
        post["th_atm"] = main["th"].copy()
        post["precip_liq_surf_flux"] = main["precip_liq_surf"].copy()
        post["precip_ice_surf_flux"] = main["precip_ice_surf"].copy()
        return run_post(post), main

    def o_pre(f):
        oracle.preamble(f); return f

    def o_main(m):
        oracle.p3_main(m, dt, True, True); return m

    def o_post(p):
        oracle.postamble(p, dt); return p

    h = p3lib.P3Handle(0, 8)

    def g_pre(f):
        d = _to_dev(torch, f); h.preamble(d); return {k: v.cpu().numpy() for k, v in d.items()}

    def g_main(m):
        d = p3lib.to_device(m); h.main(d, dt, True, True); return p3lib.to_host(d)

    def g_post(p):
        d = _to_dev(torch, p); h.postamble(d, dt); return {k: v.cpu().numpy() for k, v in d.items()}

    ref, ref_main = sequence(o_pre, o_main, o_post, lambda f: f)
    got, got_main = sequence(g_pre, g_main, g_post, lambda f: f)
    names = [n for n in POSTAMBLE_OUTPUTS if n in ref]
    errs = parity.field_errors(ref, got, names=names)
    # p3_main in the middle: some of the 96 random columns are ill-conditioned (see tests/parity.py), so the whole
    # sequence is compared with the multi-step tolerance class
    assert max(errs.values()) <= 1e-9, {k: v for k, v in errs.items() if v > 1e-9}
    h.close()


def _fused_case(torch, p3lib, ic, ncol, nlev, dtype, seed=11):
    """Device state after the preamble: p3_main fields + the extra postamble arrays; the postamble dict ALIASES the
    p3_main tensors it shares with them."""
    from scream_b200 import ic_cases
    from scream_b200.fields import FIELDS
    base = (ic_cases.make_mixed_tiled(ncol, nlev, tile=3, dtype=dtype) if ic == "mixed"
            else ic_cases.make_random(ncol, nlev, dtype=dtype))
    parity.zero_outputs(base)
    extra = prepost_cases.make_postamble(ncol, nlev, dtype, seed=seed)
    main = {n: torch.from_numpy(np.ascontiguousarray(base[n], dtype=dtype)).cuda() for n in FIELDS}
    post = {n: torch.from_numpy(extra[n]).cuda() for n in ("pseudo_density_dry", "T_prev", "qv_prev", "precip_liq_surf_mass",
                                                            "precip_ice_surf_mass", "vapor_flux", "water_flux", "ice_flux",
                                                            "heat_flux")}
    post["pseudo_density_dry"] = (main["dpres"] * (1 - main["qv"])).contiguous()
    post["T_atm"] = (main["th"] / main["inv_exner"]).contiguous() + 0.25
    post.update(pmid=main["pres"], pseudo_density=main["dpres"], th_atm=main["th"],
                precip_liq_surf_flux=main["precip_liq_surf"], precip_ice_surf_flux=main["precip_ice_surf"])
    for n in ("qv", "qc", "nc", "qr", "nr", "qi", "qm", "ni", "bm", "diag_eff_radius_qc", "diag_eff_radius_qi",
              "diag_eff_radius_qr"):
        post[n] = main[n]
    return main, post


@pytest.mark.parametrize("rb,ic,ncol,nlev", [(8, "mixed", 384, 128), (8, "random", 517, 72), (4, "random", 384, 128)])
def test_fused_main_postamble_is_bit_identical_to_the_two_calls(p3lib, rb, ic, ncol, nlev):
    """p3b200_main_postamble (postamble as the tail of phase C) == p3b200_main then p3b200_postamble_checked,
    bit for bit in every array either of them touches, and with the same 15 check results."""
    torch = pytest.importorskip("torch")
    dtype = np.float64 if rb == 8 else np.float32
    h = p3lib.P3Handle(0, rb)
    m1, p1 = _fused_case(torch, p3lib, ic, ncol, nlev, dtype)
    h.main(m1, 300.0, True, True)
    c1 = h.postamble_checked(p1, 300.0)
    d1 = h.get_diag(ncol)
    m2, p2 = _fused_case(torch, p3lib, ic, ncol, nlev, dtype)
    c2 = h.main_postamble(m2, p2, 300.0, True, True, checks=True)
    d2 = h.get_diag(ncol)
    assert np.array_equal(d1, d2)
    for name, a, b in [(n, m1[n], m2[n]) for n in m1] + [(n, p1[n], p2[n]) for n in p1]:
        assert np.array_equal(a.cpu().numpy(), b.cpu().numpy(), equal_nan=True), name
    for n in c1:
        for k in c1[n]:
            v1, v2 = c1[n][k], c2[n][k]
            assert v1 == v2 or (v1 != v1 and v2 != v2), (n, k, v1, v2)
    # without the checks the call stays asynchronous and writes the same arrays
    m3, p3 = _fused_case(torch, p3lib, ic, ncol, nlev, dtype)
    assert h.main_postamble(m3, p3, 300.0, True, True) is None
    torch.cuda.synchronize()
    for n in p1:
        assert np.array_equal(p1[n].cpu().numpy(), p3[n].cpu().numpy(), equal_nan=True), n
    h.close()


def test_fused_main_postamble_rejects_foreign_arrays(p3lib):
    torch = pytest.importorskip("torch")
    h = p3lib.P3Handle(0, 8)
    m, p = _fused_case(torch, p3lib, "mixed", 96, 72, np.float64)
    p = dict(p)
    p["qc"] = p["qc"].clone()          # not the array p3_main writes
    with pytest.raises(Exception):
        h.main_postamble(m, p, 300.0, True, True)
    h.close()


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
@pytest.mark.parametrize("ic", ["mixed", "random"])
def test_preamble_fused_into_p3_main_is_bit_identical_to_the_two_calls(p3lib, dtype, ic):
    """p3b200_main_preamble (the p3_preamble of every level run inside the streaming classification pass) against
    p3b200_preamble followed by p3b200_main on the same field-manager state: every array either call writes."""
    torch = pytest.importorskip("torch")
    from scream_b200.fields import FIELDS
    from tests import test_gpu_process as tp
    ncol, nlev, dt = 700, 128, 300.0
    fm = tp.field_manager_state(ncol, nlev, seed=3, ic=ic)
    rb = np.dtype(dtype).itemsize
    h = p3lib.P3Handle(0, rb)

    def run(fused):
        d = {k: torch.from_numpy(v.astype(dtype)).cuda() for k, v in fm.items()}
        new = lambda *s: torch.full(s, -7.0, dtype=d["qv"].dtype, device="cuda")  # noqa: E731
        buf = {n: new(ncol, nlev) for n in ("inv_exner", "th_atm", "cld_frac_l", "cld_frac_i", "dz")}
        pre = {"pmid": d["p_mid"], "pseudo_density": d["pseudo_density"], "pseudo_density_dry": d["pseudo_density_dry"],
               "T_atm": d["T_mid"], "cld_frac_t": d["cldfrac_tot"], "qv_prev": d["qv_prev_micro_step"],
               "cld_frac_r": d["rainfrac"], **buf}
        for n in ("qv", "qc", "nc", "qr", "nr", "qi", "qm", "ni", "bm"):
            pre[n] = d[n]
        main = {"qc": d["qc"], "nc": d["nc"], "qr": d["qr"], "nr": d["nr"], "qi": d["qi"], "qm": d["qm"], "ni": d["ni"],
                "bm": d["bm"], "qv": d["qv"], "th": buf["th_atm"], "nc_nuceat_tend": d["nc_nuceat_tend"], "nccn": d["nccn"],
                "ni_activated": d["ni_activated"], "inv_qc_relvar": d["inv_qc_relvar"], "cld_frac_i": buf["cld_frac_i"],
                "cld_frac_l": buf["cld_frac_l"], "cld_frac_r": d["rainfrac"], "pres": d["p_dry_mid"], "dz": buf["dz"],
                "dpres": d["pseudo_density_dry"], "inv_exner": buf["inv_exner"], "qv_prev": d["qv_prev_micro_step"],
                "t_prev": d["T_prev_micro_step"], "qv2qi_depos_tend": new(ncol, nlev), "diag_eff_radius_qc": d["eff_radius_qc"],
                "diag_eff_radius_qi": d["eff_radius_qi"], "diag_eff_radius_qr": d["eff_radius_qr"], "rho_qi": new(ncol, nlev),
                "precip_liq_flux": new(ncol, nlev + 1), "precip_ice_flux": new(ncol, nlev + 1),
                "precip_total_tend": d["precip_total_tend"], "nevapr": d["nevapr"],
                "liq_ice_exchange": d["micro_liq_ice_exchange"], "vap_liq_exchange": d["micro_vap_liq_exchange"],
                "vap_ice_exchange": d["micro_vap_ice_exchange"], "precip_liq_surf": new(ncol), "precip_ice_surf": new(ncol)}
        assert set(main) == set(FIELDS)
        if fused:
            h.main_preamble(main, pre, dt, True, True, zero_history=True)
        else:
            h.preamble(pre)
            h.main(main, dt, True, True, zero_history=True)
        torch.cuda.synchronize()
        out = {k: v.cpu().numpy() for k, v in main.items()}
        out.update({"pre_" + k: v.cpu().numpy() for k, v in pre.items()})
        return out, h.get_diag(ncol)

    (a, da), (b, db) = run(False), run(True)
    assert np.array_equal(da, db)
    for n in a:
        assert np.array_equal(a[n], b[n], equal_nan=True), n
    h.close()
