"""Field order of the p3_main argument structs as flattened by the C ABI (include/p3_b200.h).

Mirrors P3PrognosticState / P3DiagnosticInputs / P3DiagnosticOutputs / P3HistoryOnly
(reference: components/eamxx/src/physics/p3/p3_functions.hpp:161-276).
"""

PROGNOSTIC = ["qc", "nc", "qr", "nr", "qi", "qm", "ni", "bm", "qv", "th"]
DIAG_INPUTS = ["nc_nuceat_tend", "nccn", "ni_activated", "inv_qc_relvar", "cld_frac_i", "cld_frac_l",
               "cld_frac_r", "pres", "dz", "dpres", "inv_exner", "qv_prev", "t_prev"]
DIAG_OUTPUTS_2D = ["qv2qi_depos_tend", "diag_eff_radius_qc", "diag_eff_radius_qi", "diag_eff_radius_qr",
                   "rho_qi", "precip_liq_flux", "precip_ice_flux", "precip_total_tend", "nevapr"]
HISTORY = ["liq_ice_exchange", "vap_liq_exchange", "vap_ice_exchange"]
SURFACE_1D = ["precip_liq_surf", "precip_ice_surf"]

FIELDS = PROGNOSTIC + DIAG_INPUTS + DIAG_OUTPUTS_2D + HISTORY + SURFACE_1D
NUM_FIELDS = len(FIELDS)            # 37
INDEX = {n: i for i, n in enumerate(FIELDS)}

# fields with nlev+1 levels (p3_functions.hpp:232-235, "pverp")
INTERFACE_FIELDS = ("precip_liq_flux", "precip_ice_flux")
OUTPUTS = PROGNOSTIC + DIAG_OUTPUTS_2D + HISTORY + SURFACE_1D

# integer diagnostics exported per column (bit-exact parity targets)
DIAG_NAMES = ["nucleationPossible", "hydrometeorsPresent_part1", "hydrometeorsPresent_part2",
              "qc_ktop", "qc_kbot", "qc_nsub", "qr_ktop", "qr_kbot", "qr_nsub",
              "qi_ktop", "qi_kbot", "qi_nsub", "error", "pad13", "pad14", "pad15"]
NUM_DIAG = 16

# P3Runtime (p3_functions.hpp:112-133): order of the 20 values passed through the C ABI
RUNTIME_DEFAULTS = [
    ("max_total_ni", 740.0e3), ("autoconversion_prefactor", 1350.0), ("autoconversion_qc_exponent", 2.47),
    ("autoconversion_nc_exponent", 1.79), ("autoconversion_radius", 25.0e-6), ("accretion_prefactor", 67.0),
    ("accretion_qc_exponent", 1.15), ("accretion_qr_exponent", 1.15), ("rain_selfcollection_prefactor", 5.78),
    ("rain_selfcollection_breakup_diameter", 0.00028), ("constant_mu_rain", 1.0), ("spa_ccn_to_nc_factor", 1.0),
    ("cldliq_to_ice_collection_factor", 0.5), ("rain_to_ice_collection_factor", 1.0), ("min_rime_rho", 50.0),
    ("max_rime_rho", 900.0), ("immersion_freezing_exponent", 0.65), ("deposition_nucleation_exponent", 0.304),
    ("ice_sedimentation_factor", 1.0), ("do_ice_production", 1.0),
]


def levels_of(name, nlev):
    if name in SURFACE_1D:
        return 1
    return nlev + 1 if name in INTERFACE_FIELDS else nlev


def algorithmic_bytes_per_col_level(real_bytes, predict_nc=True, prescribed_ccn=True):
    """SURVEY.md §8(d) / BASELINE.md §3: every live interface array touched once."""
    reads = 19 + (1 if prescribed_ccn else 0) + (2 if (predict_nc and not prescribed_ccn) else 0)
    return (reads + 22) * real_bytes


# ---- P3Microphysics::run_impl brackets (eamxx_p3_process_interface.hpp:62-262) ------------------------------
# Names in the order of the pointer tables of oracle/p3_oracle_prepost.cpp; the C ABI structs
# (include/p3_b200.h: P3B200PreambleFields / P3B200PostambleFields) use the same member names.
PREAMBLE_FIELDS = ("pmid", "pseudo_density", "pseudo_density_dry", "T_atm", "cld_frac_t",
                   "qv", "qc", "nc", "qr", "nr", "qi", "qm", "ni", "bm", "qv_prev",
                   "inv_exner", "th_atm", "cld_frac_l", "cld_frac_i", "cld_frac_r", "dz")
PREAMBLE_OUTPUTS = PREAMBLE_FIELDS[5:]
POSTAMBLE_FIELDS_2D = ("pmid", "pseudo_density", "pseudo_density_dry", "th_atm", "T_atm",
                       "qv", "qc", "nc", "qr", "nr", "qi", "qm", "ni", "bm",
                       "diag_eff_radius_qc", "diag_eff_radius_qi", "diag_eff_radius_qr", "T_prev", "qv_prev")
POSTAMBLE_FIELDS_1D = ("precip_liq_surf_flux", "precip_ice_surf_flux", "precip_liq_surf_mass", "precip_ice_surf_mass",
                       "vapor_flux", "water_flux", "ice_flux", "heat_flux")
POSTAMBLE_OUTPUTS = POSTAMBLE_FIELDS_2D[4:] + POSTAMBLE_FIELDS_1D[2:]
# compulsory HBM traffic per column-level (arrays read + written once), SURVEY.md §8(f) N1
PREAMBLE_ARRAYS_RW = (15, 16)
POSTAMBLE_ARRAYS_RW = (17, 15)

# The property checks P3Microphysics registers (eamxx_p3_process_interface.cpp:226-243), in registration order:
# (check field, lower bound, upper bound, can_repair).  Names on the left are the C-ABI members of
# P3B200CheckFields; CHECK_POSTAMBLE_SOURCE maps each to the postamble field that holds it.
CHECKS = (
    ("T_mid", 100.0, 500.0, False), ("qv", 1e-13, 0.2, True),
    ("qc", 0.0, 0.1, False), ("qi", 0.0, 0.1, False), ("qr", 0.0, 0.1, False), ("qm", 0.0, 0.1, False),
    ("nc", 0.0, 1.0e11, False), ("nr", 0.0, 1.0e10, False), ("ni", 0.0, 1.0e10, False), ("bm", 0.0, 1.0, False),
    ("precip_liq_surf_mass", 0.0, float.fromhex("0x1.fffffffffffffp+1023"), False),
    ("precip_ice_surf_mass", 0.0, float.fromhex("0x1.fffffffffffffp+1023"), False),
    ("eff_radius_qc", 0.0, 1.0e2, False), ("eff_radius_qi", 0.0, 5.0e3, False), ("eff_radius_qr", 0.0, 5.0e3, False),
)
CHECK_FIELDS = tuple(c[0] for c in CHECKS)
CHECK_COLUMN_FIELDS = ("precip_liq_surf_mass", "precip_ice_surf_mass")
CHECK_POSTAMBLE_SOURCE = {"T_mid": "T_atm", "eff_radius_qc": "diag_eff_radius_qc", "eff_radius_qi": "diag_eff_radius_qi",
                          "eff_radius_qr": "diag_eff_radius_qr"}
CHECK_PASS, CHECK_REPAIRABLE, CHECK_FAIL = 0, 1, 2

# MassAndEnergyColumnConservationCheck (share/property_checks/mass_and_energy_column_conservation_check.cpp):
# pointer members of P3B200ConservationFields in declaration order
CONSERVATION_FIELDS = ("pseudo_density", "T_mid", "horiz_winds", "qv", "qc", "qi", "qr", "ps", "phis",
                       "vapor_flux", "water_flux", "ice_flux", "heat_flux")
