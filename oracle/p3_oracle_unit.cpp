// ORACLE — TEST INFRASTRUCTURE ONLY (see p3_oracle_types.hpp header).
// Per-function entry points of the restatement, the analogue of the reference's *_f bridges
// (p3/p3_functions_f90.hpp:881-944), used by tests/test_cpu_oracle_properties.py to run the
// reference's own closed-form property tests (p3/tests/*_unit_tests.cpp) against the oracle.
#include "p3_oracle_main.hpp"

#include <cstring>

using namespace p3o;
using PD = Procs<double>;
using MD = Main<double>;

extern "C" {

// calc_first_order_upwind_step<2> (p3_upwind_impl.hpp:18-91): arrays of nk doubles; flux/V/r are [2][nk]
void p3o_unit_upwind2(int nk, int k_bot, int k_top, int kdir, double dt_sub, const double* rho, const double* inv_rho,
                      const double* inv_dz, double* flux, double* V, double* r) {
  double* fl[2] = {flux, flux + nk};
  double* vs[2] = {V, V + nk};
  double* rs[2] = {r, r + nk};
  MD::calc_first_order_upwind_step<2>(rho, inv_rho, inv_dz, nk, k_bot, k_top, kdir, dt_sub, fl, vs, rs);
}

// generalized_sedimentation<1> (p3_upwind_impl.hpp:97-129); returns the updated k_qxbot
int p3o_unit_generalized_sedimentation1(int nk, int k_qxtop, int k_qxbot, int kbot, int kdir, double Co_max,
                                        double* dt_left, double* prt_accum, const double* rho, const double* inv_rho,
                                        const double* inv_dz, double* flux, double* V, double* r) {
  double* fl[1] = {flux};
  double* vs[1] = {V};
  double* rs[1] = {r};
  int err = 0;
  MD::generalized_sedimentation<1>(rho, inv_rho, inv_dz, nk, k_qxtop, k_qxbot, kbot, kdir, Co_max, *dt_left, *prt_accum,
                                   fl, vs, rs, &err);
  return k_qxbot;
}

// find_top / find_bottom (p3_find_impl.hpp:57-92, 17-52); returns k, *present = found
int p3o_unit_find_top(const double* v, double small, int kbot, int ktop, int kdir, int* present) {
  bool p = false;
  const int k = MD::find_top(v, small, kbot, ktop, kdir, p);
  *present = p;
  return k;
}
int p3o_unit_find_bottom(const double* v, double small, int kbot, int ktop, int kdir, int* present) {
  bool p = false;
  const int k = MD::find_bottom(v, small, kbot, ktop, kdir, p);
  *present = p;
  return k;
}

// The 30 tendencies of PD::Tend as a flat array, in declaration order (p3_oracle_procs.hpp: struct Tend).
static_assert(sizeof(PD::Tend) == 30 * sizeof(double), "Tend is 30 doubles");
void p3o_unit_cloud_water_conservation(double qc, double dt, double* tend30) {
  PD::Tend t; std::memcpy(static_cast<void*>(&t), tend30, sizeof t);
  PD::cloud_water_conservation(qc, dt, t);
  std::memcpy(tend30, &t, sizeof t);
}
void p3o_unit_rain_water_conservation(double qr, double dt, double* tend30) {
  PD::Tend t; std::memcpy(static_cast<void*>(&t), tend30, sizeof t);
  PD::rain_water_conservation(qr, dt, t);
  std::memcpy(tend30, &t, sizeof t);
}
void p3o_unit_ice_water_conservation(double qi, double dt, double* tend30) {
  PD::Tend t; std::memcpy(static_cast<void*>(&t), tend30, sizeof t);
  PD::ice_water_conservation(qi, dt, t);
  std::memcpy(tend30, &t, sizeof t);
}

// evaporate_rain (p3_evaporate_rain_impl.hpp:67-205); in[18] in the argument order of the function
void p3o_unit_evaporate_rain(const double* in, double* qr2qv_evap_tend, double* nr_evap_tend) {
  PD::evaporate_rain(in[0], in[1], in[2], in[3], in[4], in[5], in[6], in[7], in[8], in[9], in[10], in[11], in[12],
                     in[13], in[14], in[15], in[16], in[17], *qr2qv_evap_tend, *nr_evap_tend);
}
double p3o_unit_rain_evap_tscale_weight(double x) { return PD::rain_evap_tscale_weight(x); }

// qv_sat_dry / MurphyKoop (physics_saturation_impl.hpp:41-83, 130-155)
double p3o_unit_qv_sat_dry(double t, double p, int ice) { int err = 0; return PD::qv_sat_dry(t, p, ice != 0, &err); }

// lookup_ice + apply_table_ice on a caller-supplied ice table [5][4][50][12]
// (p3_table_ice_impl.hpp:83-138, 178-217)
double p3o_unit_apply_table_ice(const double* ice_table, int idx, double qi, double ni, double qm, double rhop) {
  static P3LookupTables<double> t;
  std::memcpy(&t.ice_table_vals[0][0][0][0], ice_table, sizeof t.ice_table_vals);
  TableIce<double> tab;
  PD::lookup_ice(qi, ni, qm, rhop, tab);
  return PD::apply_table_ice(idx, t, tab);
}

}  // extern "C"
