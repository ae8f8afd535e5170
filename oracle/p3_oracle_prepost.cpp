// ORACLE — TEST INFRASTRUCTURE ONLY (see p3_oracle_types.hpp header).  Never linked into the product.
//
// Scalar restatement of the two functors that bracket p3_main in P3Microphysics::run_impl
// (components/eamxx/src/physics/p3/eamxx_p3_run.cpp:11-16, 39-44):
//   p3_preamble   eamxx_p3_process_interface.hpp:62-129
//   p3_postamble  eamxx_p3_process_interface.hpp:200-262
// with the helper formulas of share/util/scream_common_physics_functions_impl.hpp restated one to one
// (exner_function :88-97, calculate_theta_from_T :115-118, calculate_T_from_theta :162-165,
// calculate_virtual_temperature :214-222, calculate_wetmmr_from_drymmr_dp_based :327-331,
// calculate_drymmr_from_wetmmr_dp_based :375-379, calculate_dz :400-409), pack size 1, same statement order
// per level, columns and levels in the reference's loop order.  Parity unpinned like the rest of the oracle:
// the reference stores no expected outputs for these functors.
#include "p3_oracle_types.hpp"
#include "p3_oracle_math.hpp"

#include <cmath>
#include <cstdint>

namespace {

template <typename Real>
struct PF {
  using C = p3o::Constants<Real>;
  static Real exner_function(Real pressure) { return p3o::pm::pow(pressure / C::P0, C::RD * C::INV_CP); }
  static Real calculate_theta_from_T(Real temperature, Real pressure) { return temperature / exner_function(pressure); }
  static Real calculate_T_from_theta(Real theta, Real pressure) { return theta * exner_function(pressure); }
  static Real calculate_virtual_temperature(Real temperature, Real qv) {
    const Real one = 1;
    const Real c1 = -one + one / C::ep_2;
    return temperature * (one + c1 * qv);
  }
  static Real calculate_wetmmr_from_drymmr_dp_based(Real drymmr, Real pseudo_density, Real pseudo_density_dry) {
    return drymmr * pseudo_density_dry / pseudo_density;
  }
  static Real calculate_drymmr_from_wetmmr_dp_based(Real wetmmr, Real pseudo_density, Real pseudo_density_dry) {
    return wetmmr * pseudo_density / pseudo_density_dry;
  }
  static Real calculate_dz(Real pseudo_density, Real p_mid, Real T_mid, Real qv) {
    const Real T_virtual = calculate_virtual_temperature(T_mid, qv);
    return (C::RD / C::gravit) * pseudo_density * T_virtual / p_mid;
  }
};

// field order of the two pointer tables (shared with tests/oracle_lib.py)
enum Pre { PRE_PMID, PRE_PD, PRE_PDD, PRE_T_ATM, PRE_CLD_FRAC_T, PRE_QV, PRE_QC, PRE_NC, PRE_QR, PRE_NR, PRE_QI, PRE_QM,
           PRE_NI, PRE_BM, PRE_QV_PREV, PRE_INV_EXNER, PRE_TH_ATM, PRE_CLD_FRAC_L, PRE_CLD_FRAC_I, PRE_CLD_FRAC_R,
           PRE_DZ, PRE_NUM };
enum Post { POST_PMID, POST_PD, POST_PDD, POST_TH_ATM, POST_T_ATM, POST_QV, POST_QC, POST_NC, POST_QR, POST_NR, POST_QI,
            POST_QM, POST_NI, POST_BM, POST_RAD_QC, POST_RAD_QI, POST_RAD_QR, POST_T_PREV, POST_QV_PREV, POST_NUM2D,
            POST_LIQ_FLUX = POST_NUM2D, POST_ICE_FLUX, POST_LIQ_MASS, POST_ICE_MASS, POST_VAPOR_FLUX, POST_WATER_FLUX,
            POST_ICE_FLUX_OUT, POST_HEAT_FLUX, POST_NUM };

template <typename Real>
void preamble(void* const* f, int ncol, int nlev) {
  auto at = [&](int id, int i, int k) -> Real& { return static_cast<Real*>(f[id])[(int64_t)i * nlev + k]; };
  const Real mincld = Real(0.0001);
  for (int icol = 0; icol < ncol; ++icol) {
    for (int k = 0; k < nlev; ++k) {
      const Real pmid = at(PRE_PMID, icol, k), T_atm = at(PRE_T_ATM, icol, k), cld_frac_t = at(PRE_CLD_FRAC_T, icol, k);
      const Real pd = at(PRE_PD, icol, k), pdd = at(PRE_PDD, icol, k);
      at(PRE_DZ, icol, k) = PF<Real>::calculate_dz(pd, pmid, T_atm, at(PRE_QV, icol, k));
      for (int id : {PRE_QC, PRE_NC, PRE_QR, PRE_NR, PRE_QI, PRE_NI, PRE_QM, PRE_BM, PRE_QV, PRE_QV_PREV})
        at(id, icol, k) = PF<Real>::calculate_drymmr_from_wetmmr_dp_based(at(id, icol, k), pd, pdd);
      const Real exner = PF<Real>::exner_function(pmid);
      at(PRE_INV_EXNER, icol, k) = Real(1.0) / exner;
      at(PRE_TH_ATM, icol, k) = PF<Real>::calculate_theta_from_T(T_atm, pmid);
      const Real cf = cld_frac_t < mincld ? mincld : cld_frac_t;     // ekat::max(cld_frac_t, mincld)
      at(PRE_CLD_FRAC_L, icol, k) = cf;
      at(PRE_CLD_FRAC_I, icol, k) = cf;
      at(PRE_CLD_FRAC_R, icol, k) = cf;
      if (k != 0) {   // max-overlap with the layer above (l.115-129)
        const Real up = at(PRE_CLD_FRAC_T, icol, k - 1);
        at(PRE_CLD_FRAC_R, icol, k) = up > at(PRE_CLD_FRAC_R, icol, k) ? up : at(PRE_CLD_FRAC_R, icol, k);
      }
    }
  }
}

template <typename Real>
void postamble(void* const* f, int ncol, int nlev, double m_dt) {
  using C = p3o::Constants<Real>;
  auto at = [&](int id, int i, int k) -> Real& { return static_cast<Real*>(f[id])[(int64_t)i * nlev + k]; };
  auto col = [&](int id, int i) -> Real& { return static_cast<Real*>(f[id])[i]; };
  for (int icol = 0; icol < ncol; ++icol) {
    for (int k = 0; k < nlev; ++k) {
      const Real pd = at(POST_PD, icol, k), pdd = at(POST_PDD, icol, k);
      {
        const Real T_atm_before_p3 = at(POST_T_ATM, icol, k);
        at(POST_T_ATM, icol, k) = (PF<Real>::calculate_T_from_theta(at(POST_TH_ATM, icol, k), at(POST_PMID, icol, k)) - T_atm_before_p3) * pdd / pd;
        at(POST_T_ATM, icol, k) += T_atm_before_p3;
      }
      at(POST_T_PREV, icol, k) = at(POST_T_ATM, icol, k);
      for (int id : {POST_QC, POST_NC, POST_QR, POST_NR, POST_QI, POST_NI, POST_QM, POST_BM, POST_QV})
        at(id, icol, k) = PF<Real>::calculate_wetmmr_from_drymmr_dp_based(at(id, icol, k), pd, pdd);
      at(POST_QV_PREV, icol, k) = at(POST_QV, icol, k);
      at(POST_RAD_QC, icol, k) *= 1e6;
      at(POST_RAD_QI, icol, k) *= 1e6;
      at(POST_RAD_QR, icol, k) *= 1e6;
    }
    col(POST_LIQ_MASS, icol) += col(POST_LIQ_FLUX, icol) * C::RHO_H2O * m_dt;
    col(POST_ICE_MASS, icol) += col(POST_ICE_FLUX, icol) * C::RHO_H2O * m_dt;
    if (f[POST_WATER_FLUX]) {
      col(POST_VAPOR_FLUX, icol) = 0.0;
      col(POST_WATER_FLUX, icol) = col(POST_LIQ_FLUX, icol) + col(POST_ICE_FLUX, icol);
      col(POST_ICE_FLUX_OUT, icol) = col(POST_ICE_FLUX, icol);
      col(POST_HEAT_FLUX, icol) = 0.0;
    }
  }
}

}  // namespace

extern "C" {
int p3o_preamble(int real_bytes, void* const* fields, int ncol, int nlev) {
  if (real_bytes == 8) preamble<double>(fields, ncol, nlev);
  else if (real_bytes == 4) preamble<float>(fields, ncol, nlev);
  else return -1;
  return 0;
}
int p3o_postamble(int real_bytes, void* const* fields, int ncol, int nlev, double dt) {
  if (real_bytes == 8) postamble<double>(fields, ncol, nlev, dt);
  else if (real_bytes == 4) postamble<float>(fields, ncol, nlev, dt);
  else return -1;
  return 0;
}
}
