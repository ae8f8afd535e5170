// ORACLE — TEST INFRASTRUCTURE ONLY.  Not part of the product.
//
// Scalar (pack-size-1) CPU restatement of EAMxx/SCREAM P3 `p3_main`, used as the parity
// checker for the sm_100a CUDA path and as the timed "CPU port" baseline in bench.py.
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
// load this code.  The product library (scream_b200/csrc) never includes or links it.
//
// PARITY STATUS: "parity unpinned" at the p3_main level — the reference ships no stored golden
// outputs for p3_main and neither its Kokkos C++ build nor its Fortran twin can be built in this
// container (Kokkos sources missing from the archive, no gfortran).  The restatement is pinned
// by construction (same statement order as P3/impl/*.hpp, file:line cited on each function) and
// by the reference's own enabled property tests (tests/test_oracle_properties.py).
//
// All file:line citations are into /root/reference/components/eamxx/src/physics unless stated.
#ifndef P3_ORACLE_TYPES_HPP
#define P3_ORACLE_TYPES_HPP

#include <cmath>
#include <cstdint>
#include <limits>
#include <algorithm>
#include "p3_oracle_math.hpp"

namespace p3o {

// share/physics_constants.hpp:22-123 (values restated, only those P3 uses)
template <typename Real>
struct Constants {
  static constexpr Real Cpair       = 1004.64;
  static constexpr Real Rair        = 287.042;
  static constexpr Real RH2O        = 461.505;
  static constexpr Real RV          = RH2O;
  static constexpr Real RHO_H2O     = 1000.0;
  static constexpr Real INV_RHO_H2O = 1.0/RHO_H2O;
  static constexpr Real MWH2O       = 18.016;
  static constexpr Real MWdry       = 28.966;
  static constexpr Real ep_2        = MWH2O/MWdry;
  static constexpr Real gravit      = 9.80616;
  static constexpr Real LatVap      = 2501000.0;
  static constexpr Real LatIce      = 333700.0;
  static constexpr Real CpLiq       = 4188.0;
  static constexpr Real Tmelt       = 273.15;
  static constexpr Real T_zerodegc  = Tmelt;
  static constexpr Real T_homogfrz  = Tmelt - 40;
  static constexpr Real T_rainfrz   = Tmelt - 4;
  static constexpr Real Pi          = 3.14159265358979323;
  static constexpr Real RHOW        = RHO_H2O;
  static constexpr Real RHO_RIMEMIN = 50.0;
  static constexpr Real RHO_RIMEMAX = 900.0;
  static constexpr Real INV_RHO_RIMEMAX = 1.0/RHO_RIMEMAX;
  static constexpr Real THIRD       = 1.0/3.0;
  static constexpr Real SXTH        = 1.0/6.0;
  static constexpr Real PIOV3       = Pi*THIRD;
  static constexpr Real PIOV6       = Pi*SXTH;
  static constexpr Real BIMM        = 2.0;
  static constexpr Real CONS1       = PIOV6*RHOW;
  static constexpr Real CONS2       = 4.*PIOV3*RHOW;
  static constexpr Real CONS3       = 1.0/(CONS2*1.562500000000000e-14);
  static constexpr Real CONS5       = PIOV6*BIMM;
  static constexpr Real CONS6       = PIOV6*PIOV6*RHOW*BIMM;
  static constexpr Real QSMALL      = 1.e-14;
  static constexpr Real QTENDSMALL  = 1e-20;
  static constexpr Real BSMALL      = 1.e-15;
  static constexpr Real NSMALL      = 1.e-16;
  static constexpr Real ZERO        = 0.0;
  static constexpr Real ONE         = 1.0;
  static constexpr Real P0          = 100000.0;
  static constexpr Real RD          = Rair;
  static constexpr Real RHO_1000MB  = P0/(RD*Tmelt);
  static constexpr Real RHO_600MB   = 60000/(RD*253.15);
  static constexpr Real CP          = Cpair;
  static constexpr Real INV_CP      = 1.0/CP;
  static constexpr Real dt_left_tol = 1.e-4;
  static constexpr Real bcn         = 2.;
  static constexpr Real dropmass    = 5.2e-7;
  static constexpr Real NCCNST      = 200.0e+6;
  static constexpr Real incloud_limit = 5.1e-3;
  static constexpr Real precip_limit  = 1.0e-2;
  static constexpr Real f1r         = 0.78;
  static constexpr Real f2r         = 0.32;
  static constexpr Real nmltratio   = 1.0;
  static constexpr int  VTABLE_DIM0 = 300;
  static constexpr int  VTABLE_DIM1 = 10;
  static constexpr int  MU_R_TABLE_DIM = 150;
  static constexpr int  IPARAM      = 3;
};

// p3/p3_functions.hpp:30-52
struct P3C {
  enum { densize = 5, rimsize = 4, isize = 50, ice_table_size = 12, rcollsize = 30,
         collect_table_size = 2, iparam = 3, dnusize = 16 };
  static constexpr double lookup_table_1a_dum1_c = 4.135985029041767e+00;
};

// p3/p3_functions.hpp:112-133 (defaults identical)
template <typename Real>
struct P3Runtime {
  Real max_total_ni = 740.0e3;
  Real autoconversion_prefactor = 1350.0;
  Real autoconversion_qc_exponent = 2.47;
  Real autoconversion_nc_exponent = 1.79;
  Real autoconversion_radius = 25.0e-6;
  Real accretion_prefactor = 67.0;
  Real accretion_qc_exponent = 1.15;
  Real accretion_qr_exponent = 1.15;
  Real rain_selfcollection_prefactor = 5.78;
  Real rain_selfcollection_breakup_diameter = 0.00028;
  Real constant_mu_rain = 1.0;
  Real spa_ccn_to_nc_factor = 1.0;
  Real cldliq_to_ice_collection_factor = 0.5;
  Real rain_to_ice_collection_factor = 1.0;
  Real min_rime_rho = 50.0;
  Real max_rime_rho = 900.0;
  Real immersion_freezing_exponent = 0.65;
  Real deposition_nucleation_exponent = 0.304;
  Real ice_sedimentation_factor = 1.0;
  bool do_ice_production = true;
};

// p3/p3_functions.hpp:279-290 — lookup tables, row-major, in the precision of the run.
template <typename Real>
struct P3LookupTables {
  Real mu_r_table_vals[150];
  Real vn_table_vals[300][10];
  Real vm_table_vals[300][10];
  Real revap_table_vals[300][10];
  Real ice_table_vals[5][4][50][12];
  Real collect_table_vals[5][4][50][30][2];   // stored as log10 (p3_table_ice_impl.hpp:63)
  Real dnu_table_vals[16];
};

// p3/p3_functions.hpp:333-346
template <typename Real> struct Table3   { int dumii, dumjj; Real rdumii, rdumjj; };
template <typename Real> struct TableIce { int dumi, dumjj, dumii, dumzz; Real dum1, dum4, dum5, dum6; };
template <typename Real> struct TableRain{ int dumj; Real dum3; };

// One column's worth of pointers (level index fastest, level 0 = model top;
// p3_main_impl.hpp:166-204 takes the same per-column subviews).
template <typename Real>
struct Column {
  // P3PrognosticState (p3_functions.hpp:161-183)
  Real *qc, *nc, *qr, *nr, *qi, *qm, *ni, *bm, *qv, *th;
  // P3DiagnosticInputs (:186-214)
  const Real *nc_nuceat_tend, *nccn, *ni_activated, *inv_qc_relvar, *cld_frac_i, *cld_frac_l,
             *cld_frac_r, *pres, *dz, *dpres, *inv_exner, *qv_prev, *t_prev;
  // P3DiagnosticOutputs (:217-241)
  Real *qv2qi_depos_tend, *diag_eff_radius_qc, *diag_eff_radius_qi, *diag_eff_radius_qr, *rho_qi,
       *precip_liq_flux, *precip_ice_flux, *precip_total_tend, *nevapr;
  Real *precip_liq_surf, *precip_ice_surf;   // scalars of this column
  // P3HistoryOnly (:268-276)
  Real *liq_ice_exchange, *vap_liq_exchange, *vap_ice_exchange;
};

// Integer diagnostics exported per column (bit-exact parity targets).
enum DiagIdx {
  D_NUC_POSSIBLE = 0,   // nucleationPossible after part1
  D_HYDRO_PART1,        // hydrometeorsPresent after part1
  D_HYDRO_PART2,        // hydrometeorsPresent after part2 (-1 if part2 not reached)
  D_QC_KTOP, D_QC_KBOT, D_QC_NSUB,   // cloud sedimentation: k_qxtop, final k_qxbot, #substeps
  D_QR_KTOP, D_QR_KBOT, D_QR_NSUB,
  D_QI_KTOP, D_QI_KBOT, D_QI_NSUB,
  D_ERROR,              // bit 0: T<=0, bit 1: T NaN (check_temperature), bit 2: Co_max<0
  D_NUM = 16
};

} // namespace p3o
#endif
