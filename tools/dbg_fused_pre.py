"""Development helper: one fused p3b200_main_preamble on a tiled mixed field-manager state (for ncu)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from scream_b200 import lib as p3
from scream_b200.fields import FIELDS
from tests import test_gpu_process as tp
ncol, nlev, dt = int(sys.argv[1]) if len(sys.argv) > 1 else 148000, 128, 300.0
fm = tp.field_manager_state(48, nlev, ic="mixed")
idx = torch.arange(ncol, device="cuda") % 48
d = {k: torch.from_numpy(v).cuda().index_select(0, idx).contiguous() for k, v in fm.items()}
new = lambda *s: torch.full(s, -7.0, dtype=torch.float64, device="cuda")
buf = {n: new(ncol, nlev) for n in ("inv_exner", "th_atm", "cld_frac_l", "cld_frac_i", "dz")}
pre = {"pmid": d["p_mid"], "pseudo_density": d["pseudo_density"], "pseudo_density_dry": d["pseudo_density_dry"], "T_atm": d["T_mid"],
       "cld_frac_t": d["cldfrac_tot"], "qv_prev": d["qv_prev_micro_step"], "cld_frac_r": d["rainfrac"], **buf}
for n in ("qv", "qc", "nc", "qr", "nr", "qi", "qm", "ni", "bm"):
    pre[n] = d[n]
main = {"qc": d["qc"], "nc": d["nc"], "qr": d["qr"], "nr": d["nr"], "qi": d["qi"], "qm": d["qm"], "ni": d["ni"], "bm": d["bm"], "qv": d["qv"],
        "th": buf["th_atm"], "nc_nuceat_tend": d["nc_nuceat_tend"], "nccn": d["nccn"], "ni_activated": d["ni_activated"],
        "inv_qc_relvar": d["inv_qc_relvar"], "cld_frac_i": buf["cld_frac_i"], "cld_frac_l": buf["cld_frac_l"], "cld_frac_r": d["rainfrac"],
        "pres": d["p_dry_mid"], "dz": buf["dz"], "dpres": d["pseudo_density_dry"], "inv_exner": buf["inv_exner"],
        "qv_prev": d["qv_prev_micro_step"], "t_prev": d["T_prev_micro_step"], "qv2qi_depos_tend": new(ncol, nlev),
        "diag_eff_radius_qc": d["eff_radius_qc"], "diag_eff_radius_qi": d["eff_radius_qi"], "diag_eff_radius_qr": d["eff_radius_qr"],
        "rho_qi": new(ncol, nlev), "precip_liq_flux": new(ncol, nlev + 1), "precip_ice_flux": new(ncol, nlev + 1),
        "precip_total_tend": d["precip_total_tend"], "nevapr": d["nevapr"], "liq_ice_exchange": d["micro_liq_ice_exchange"],
        "vap_liq_exchange": d["micro_vap_liq_exchange"], "vap_ice_exchange": d["micro_vap_ice_exchange"],
        "precip_liq_surf": new(ncol), "precip_ice_surf": new(ncol)}
h = p3.P3Handle(0, 8)
h.main_preamble(main, pre, dt, True, True, zero_history=True)
torch.cuda.synchronize()
print(h.last_phase_ms())
