// p3_host.hpp — host-side helpers of libp3b200: lookup-table I/O and generation, and the conversion
// of the C-ABI P3B200Runtime into the device Runtime<Real> (with the host-precomputed scalars).
// Plain C++17, no CUDA: compiled into the product library and into the tests/hostsim debug build.
#ifndef P3B200_HOST_HPP
#define P3B200_HOST_HPP

#include "../../include/p3_b200.h"
#include "p3_physics.cuh"

#include <cmath>

namespace p3b200 {

// P3Runtime defaults, p3_functions.hpp:114-133
void runtime_defaults(P3B200Runtime& r);

// init_kokkos_ice_lookup_tables, p3_table_ice_impl.hpp:18-78: text table -> ice / log10(collect).
int tables_read_text(const char* path, const char* expect_version, P3B200Tables& out);
// blob = "P3TB" | u32 len | version | 12000 f64 ice | 60000 f64 collect (raw, log10 applied here)
int tables_read_blob(const char* path, const char* expect_version, P3B200Tables& out);
// p3_init_b, components/eam/src/physics/p3/scream/micro_p3.F90:218-341, + dnu p3_table3_impl.hpp:120-135
void tables_generate_rain(P3B200Tables& out);

// Device-precision runtime: every knob cast to Real the way P3Runtime's Scalar members hold them,
// plus scalars the reference recomputes per level from run-time constants.  The math is done with
// the Real overloads of <cmath> so the float build rounds like Functions<float,...> would.
template <typename Real>
Runtime<Real> make_runtime(const P3B200Runtime& r) {
  using C = Cst<Real>;
  using P = Phys<Real>;
  Runtime<Real> o;
  o.max_total_ni = Real(r.max_total_ni);
  o.autoconversion_prefactor = Real(r.autoconversion_prefactor);
  o.autoconversion_qc_exponent = Real(r.autoconversion_qc_exponent);
  o.autoconversion_nc_exponent = Real(r.autoconversion_nc_exponent);
  o.autoconversion_radius = Real(r.autoconversion_radius);
  o.accretion_prefactor = Real(r.accretion_prefactor);
  o.accretion_qc_exponent = Real(r.accretion_qc_exponent);
  o.accretion_qr_exponent = Real(r.accretion_qr_exponent);
  o.rain_selfcollection_prefactor = Real(r.rain_selfcollection_prefactor);
  o.rain_selfcollection_breakup_diameter = Real(r.rain_selfcollection_breakup_diameter);
  o.constant_mu_rain = Real(r.constant_mu_rain);
  o.spa_ccn_to_nc_factor = Real(r.spa_ccn_to_nc_factor);
  o.cldliq_to_ice_collection_factor = Real(r.cldliq_to_ice_collection_factor);
  o.rain_to_ice_collection_factor = Real(r.rain_to_ice_collection_factor);
  o.min_rime_rho = Real(r.min_rime_rho);
  o.max_rime_rho = Real(r.max_rime_rho);
  o.immersion_freezing_exponent = Real(r.immersion_freezing_exponent);
  o.deposition_nucleation_exponent = Real(r.deposition_nucleation_exponent);
  o.ice_sedimentation_factor = Real(r.ice_sedimentation_factor);
  o.do_ice_production = r.do_ice_production ? 1 : 0;
  // p3_autoconversion_impl.hpp:30-36
  if (o.autoconversion_radius == Real(25.0e-6)) o.autoconv_cons3 = C::CONS3;
  else o.autoconv_cons3 = Real(1.0) / (C::CONS2 * P::x_pow(o.autoconversion_radius, Real(3.0)));
  // MurphyKoop_svp(Tmelt, liquid): physics_saturation_impl.hpp:62-71 at T = 273.15
  const Real tm = C::Tmelt;
  o.svp_liq_tmelt = P::svp_liquid(tm, P::x_log(tm));
  const Real mu_r = o.constant_mu_rain;
  o.gamma_mur_p1 = P::x_tgamma(mu_r + 1);
  o.gamma_mur_p2 = P::x_tgamma(mu_r + 2);
  o.lgamma7_mur = P::x_log(P::x_tgamma(Real(7.) + mu_r));
  o.lgamma4_mur = P::x_log(P::x_tgamma(Real(4.) + mu_r));
  return o;
}

}  // namespace p3b200
#endif
