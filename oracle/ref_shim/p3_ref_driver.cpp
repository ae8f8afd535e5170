// p3_ref_driver.cpp — TEST INFRASTRUCTURE ONLY.
// A C entry point around the reference's OWN p3_main: this translation unit #includes the unmodified headers under
// /root/reference (components/eamxx/src/physics/p3/p3_functions.hpp, impl/*.hpp, physics/share/*.hpp and the EKAT
// headers they pull in) and compiles them against oracle/ref_shim/Kokkos_Core.hpp.  Built by oracle/Makefile.ref
// into oracle/_ref/libp3ref_<dp|sp>_pk<N>.so; used by tests/test_cpu_oracle_vs_ref.py to pin the oracle
// restatement, by tools/make_ref_golden.py to generate the committed fixtures, and (optionally) by bench.py as
// the CPU baseline of kind "reference".  Never linked into the product.
//
// Same calling convention as oracle/p3_oracle.cpp (p3o_main): host pointers, (ncol, nlev) level-contiguous arrays
// with a per-field column stride, field order of include/p3_b200.h.  What this file does itself is only what the
// reference's test bridge p3_main_f does around the call (p3_functions_f90.cpp:1888-2103): copy the caller's
// arrays into Kokkos views of packs, build the argument structs, construct the WorkspaceManager the way
// P3Microphysics::initialize_impl does (eamxx_p3_process_interface.cpp:141,424: 52 slots), call
// Functions<Real,DefaultDevice>::p3_main and copy the results back.
#include "physics/p3/p3_functions.hpp"

#include "physics/p3/impl/p3_table3_impl.hpp"
#include "physics/p3/impl/p3_table_ice_impl.hpp"
#include "physics/p3/impl/p3_back_to_cell_average_impl.hpp"
#include "physics/p3/impl/p3_dsd2_impl.hpp"
#include "physics/p3/impl/p3_upwind_impl.hpp"
#include "physics/p3/impl/p3_find_impl.hpp"
#include "physics/p3/impl/p3_conservation_impl.hpp"
#include "physics/p3/impl/p3_autoconversion_impl.hpp"
#include "physics/p3/impl/p3_rain_self_collection_impl.hpp"
#include "physics/p3/impl/p3_impose_max_total_ni_impl.hpp"
#include "physics/p3/impl/p3_calc_rime_density_impl.hpp"
#include "physics/p3/impl/p3_cldliq_imm_freezing_impl.hpp"
#include "physics/p3/impl/p3_droplet_self_coll_impl.hpp"
#include "physics/p3/impl/p3_cloud_sed_impl.hpp"
#include "physics/p3/impl/p3_cloud_rain_acc_impl.hpp"
#include "physics/p3/impl/p3_ice_sed_impl.hpp"
#include "physics/p3/impl/p3_rain_sed_impl.hpp"
#include "physics/p3/impl/p3_rain_imm_freezing_impl.hpp"
#include "physics/p3/impl/p3_get_time_space_phys_variables_impl.hpp"
#include "physics/p3/impl/p3_evaporate_rain_impl.hpp"
#include "physics/p3/impl/p3_update_prognostics_impl.hpp"
#include "physics/p3/impl/p3_ice_collection_impl.hpp"
#include "physics/p3/impl/p3_ice_deposition_sublimation_impl.hpp"
#include "physics/p3/impl/p3_ice_relaxation_timescale_impl.hpp"
#include "physics/p3/impl/p3_ice_nucleation_impl.hpp"
#include "physics/p3/impl/p3_ice_melting_impl.hpp"
#include "physics/p3/impl/p3_calc_liq_relaxation_timescale_impl.hpp"
#include "physics/p3/impl/p3_ice_cldliq_wet_growth_impl.hpp"
#include "physics/p3/impl/p3_check_values_impl.hpp"
#include "physics/p3/impl/p3_incloud_mixingratios_impl.hpp"
#include "physics/p3/impl/p3_subgrid_variance_scaling_impl.hpp"
#include "physics/p3/impl/p3_main_impl.hpp"
#include "physics/p3/impl/p3_main_impl_part1.hpp"
#include "physics/p3/impl/p3_main_impl_part2.hpp"
#include "physics/p3/impl/p3_main_impl_part3.hpp"
#include "physics/p3/impl/p3_ice_supersat_conservation_impl.hpp"
#include "physics/p3/impl/p3_nc_conservation_impl.hpp"
#include "physics/p3/impl/p3_nr_conservation_impl.hpp"
#include "physics/p3/impl/p3_ni_conservation_impl.hpp"
#include "physics/p3/impl/p3_prevent_liq_supersaturation_impl.hpp"
#include "physics/share/physics_saturation_impl.hpp"

#include <chrono>
#include <cstdint>
#include <cstring>
#include <vector>

namespace {

using scream::Real;
using scream::Int;
using P3F = scream::p3::Functions<Real, scream::DefaultDevice>;
using Spack = typename P3F::Spack;
using view_2d = typename P3F::view_2d<Spack>;
using view_1d = typename P3F::view_1d<Real>;
constexpr int PS = Spack::n;

// field order of include/p3_b200.h / oracle/p3_oracle.cpp
enum Field {
  F_QC, F_NC, F_QR, F_NR, F_QI, F_QM, F_NI, F_BM, F_QV, F_TH,
  F_NC_NUCEAT_TEND, F_NCCN, F_NI_ACTIVATED, F_INV_QC_RELVAR, F_CLD_FRAC_I, F_CLD_FRAC_L, F_CLD_FRAC_R,
  F_PRES, F_DZ, F_DPRES, F_INV_EXNER, F_QV_PREV, F_T_PREV,
  F_QV2QI_DEPOS_TEND, F_DIAG_EFF_RADIUS_QC, F_DIAG_EFF_RADIUS_QI, F_DIAG_EFF_RADIUS_QR, F_RHO_QI,
  F_PRECIP_LIQ_FLUX, F_PRECIP_ICE_FLUX, F_PRECIP_TOTAL_TEND, F_NEVAPR,
  F_LIQ_ICE_EXCHANGE, F_VAP_LIQ_EXCHANGE, F_VAP_ICE_EXCHANGE,
  F_PRECIP_LIQ_SURF, F_PRECIP_ICE_SURF, F_NUM
};

// rain tables handed to the reference's init_kokkos_tables through the Fortran hook it calls
const double* g_vn = nullptr;
const double* g_vm = nullptr;
const double* g_revap = nullptr;
const double* g_mu_r = nullptr;

struct RefTables {
  P3F::P3LookupTables t;
};

// caller array (ncol x nlev_f, column stride cs) -> view of packs (ncol x npack(nlev_f)); pad lanes get `pad`
view_2d to_view(const char* name, const Real* src, int64_t cs, int ncol, int nlev_f, Real pad) {
  const int np = (nlev_f + PS - 1) / PS;
  view_2d v(name, ncol, np);
  Real* flat = reinterpret_cast<Real*>(v.data());
  for (int i = 0; i < ncol; ++i) {
    Real* row = flat + int64_t(i) * np * PS;
    for (int k = 0; k < nlev_f; ++k) row[k] = src[int64_t(i) * cs + k];
    for (int k = nlev_f; k < np * PS; ++k) row[k] = pad;
  }
  return v;
}
void from_view(const view_2d& v, Real* dst, int64_t cs, int ncol, int nlev_f) {
  const int np = int(v.extent(1));
  const Real* flat = reinterpret_cast<const Real*>(v.data());
  for (int i = 0; i < ncol; ++i)
    for (int k = 0; k < nlev_f; ++k) dst[int64_t(i) * cs + k] = flat[int64_t(i) * np * PS + k];
}

}  // namespace

// The Fortran routine the reference's init_kokkos_tables calls (p3_iso_c.f90:26-41): column-major (300,10) arrays.
// The Fortran compiler is absent, so the values come from the caller (generated by the repo's restatement of
// micro_p3.F90:218-341, p3_init_b) — the one input of this build that is not the reference's own code.
extern "C" void scream::p3::init_tables_from_f90_c(Real* vn, Real* vm, Real* revap, Real* mu) {
  for (int i = 0; i < 300; ++i)
    for (int j = 0; j < 10; ++j) {
      vn[j * 300 + i] = Real(g_vn[i * 10 + j]);
      vm[j * 300 + i] = Real(g_vm[i * 10 + j]);
      revap[j * 300 + i] = Real(g_revap[i * 10 + j]);
    }
  for (int i = 0; i < 150; ++i) mu[i] = Real(g_mu_r[i]);
}

extern "C" {

int p3ref_real_bytes() { return int(sizeof(Real)); }
int p3ref_pack_size() { return PS; }
int p3ref_max_threads() { return Kokkos::OpenMP().concurrency(); }

// Tables from host arrays: ice[5][4][50][12], coll_raw[5][4][50][30][2] (raw values; the reference stores log10),
// vn/vm/revap[300][10] row-major, mu_r[150].  The rain tables go through the reference's own init_kokkos_tables.
void* p3ref_tables_create(const double* ice, const double* coll_raw, const double* vn, const double* vm,
                          const double* revap, const double* mu_r) {
  auto* h = new RefTables;
  g_vn = vn; g_vm = vm; g_revap = revap; g_mu_r = mu_r;
  P3F::init_kokkos_tables(h->t.vn_table_vals, h->t.vm_table_vals, h->t.revap_table_vals, h->t.mu_r_table_vals,
                          h->t.dnu_table_vals);
  g_vn = g_vm = g_revap = g_mu_r = nullptr;
  typename P3F::view_ice_table::non_const_type ice_v("ice_table_vals");
  typename P3F::view_collect_table::non_const_type coll_v("collect_table_vals");
  for (int i = 0; i < 12000; ++i) ice_v.data()[i] = Real(ice[i]);
  for (int i = 0; i < 60000; ++i) coll_v.data()[i] = Real(std::log10(coll_raw[i]));
  h->t.ice_table_vals = ice_v;
  h->t.collect_table_vals = coll_v;
  return h;
}

// Ice / collection tables through the reference's own text-file loader (p3_table_ice_impl.hpp:18-78); the path is
// the compile-time SCREAM_DATA_DIR "/tables/p3_lookup_table_1.dat-v4.1.1".  Only meaningful where /root/reference
// exists.  Copies the loaded tables out ([12000] and [60000], the latter already log10) and installs them in h.
int p3ref_tables_load_ice_with_reference_loader(void* handle, double* ice_out, double* coll_log10_out) {
  auto* h = static_cast<RefTables*>(handle);
  try {
    P3F::init_kokkos_ice_lookup_tables(h->t.ice_table_vals, h->t.collect_table_vals);
  } catch (const std::exception& e) {
    std::fprintf(stderr, "p3ref: %s\n", e.what());
    return -1;
  }
  for (int i = 0; i < 12000; ++i) ice_out[i] = double(h->t.ice_table_vals.data()[i]);
  for (int i = 0; i < 60000; ++i) coll_log10_out[i] = double(h->t.collect_table_vals.data()[i]);
  return 0;
}

void p3ref_tables_destroy(void* handle) { delete static_cast<RefTables*>(handle); }

// p3_main over ncol columns.  runtime: 20 doubles in P3Runtime member order, or NULL.  Returns the reference's own
// elapsed time (microseconds measured inside p3_main_internal, p3_main_impl.hpp:105,326-330) in seconds, or < 0.
double p3ref_main(void* tables, int real_bytes, void* const* fields, const int64_t* strides, int ncol, int nlev,
                  double dt, int predictNc, int prescribedCCN, const double* runtime, int* /*diag*/, int nthreads) {
  if (real_bytes != int(sizeof(Real))) return -2.0;
  auto* h = static_cast<RefTables*>(tables);
  if (!h) return -1.0;
#ifdef _OPENMP
  const int saved_threads = omp_get_max_threads();
  if (nthreads > 0) omp_set_num_threads(nthreads);
#endif
  auto in = [&](int f) { return static_cast<const Real*>(fields[f]); };
  auto out = [&](int f) { return static_cast<Real*>(fields[f]); };

  P3F::P3Runtime ro;
  if (runtime) {
    const double* r = runtime;
    ro.max_total_ni = r[0]; ro.autoconversion_prefactor = r[1]; ro.autoconversion_qc_exponent = r[2];
    ro.autoconversion_nc_exponent = r[3]; ro.autoconversion_radius = r[4]; ro.accretion_prefactor = r[5];
    ro.accretion_qc_exponent = r[6]; ro.accretion_qr_exponent = r[7]; ro.rain_selfcollection_prefactor = r[8];
    ro.rain_selfcollection_breakup_diameter = r[9]; ro.constant_mu_rain = r[10]; ro.spa_ccn_to_nc_factor = r[11];
    ro.cldliq_to_ice_collection_factor = r[12]; ro.rain_to_ice_collection_factor = r[13];
    ro.min_rime_rho = r[14]; ro.max_rime_rho = r[15]; ro.immersion_freezing_exponent = r[16];
    ro.deposition_nucleation_exponent = r[17]; ro.ice_sedimentation_factor = r[18];
    ro.do_ice_production = r[19] != 0.0;
  }

  // Pad lanes (pack sizes > 1 with nlev % PS != 0) hold benign values: positive thermodynamic inputs so that the
  // saturation functions' assertions hold, zero water.
  auto v = [&](const char* n, int f, Real pad, int extra = 0) { return to_view(n, in(f), strides[f], ncol, nlev + extra, pad); };

  P3F::P3PrognosticState prog;
  prog.qc = v("qc", F_QC, 0); prog.nc = v("nc", F_NC, 0); prog.qr = v("qr", F_QR, 0); prog.nr = v("nr", F_NR, 0);
  prog.qi = v("qi", F_QI, 0); prog.qm = v("qm", F_QM, 0); prog.ni = v("ni", F_NI, 0); prog.bm = v("bm", F_BM, 0);
  prog.qv = v("qv", F_QV, Real(1e-4)); prog.th = v("th", F_TH, Real(300));

  P3F::P3DiagnosticInputs din;
  din.nc_nuceat_tend = v("nc_nuceat_tend", F_NC_NUCEAT_TEND, 0);
  din.nccn = v("nccn", F_NCCN, 0);
  din.ni_activated = v("ni_activated", F_NI_ACTIVATED, 0);
  din.inv_qc_relvar = v("inv_qc_relvar", F_INV_QC_RELVAR, 1);
  din.cld_frac_i = v("cld_frac_i", F_CLD_FRAC_I, 1);
  din.cld_frac_l = v("cld_frac_l", F_CLD_FRAC_L, 1);
  din.cld_frac_r = v("cld_frac_r", F_CLD_FRAC_R, 1);
  din.pres = v("pres", F_PRES, Real(1e5)); din.dz = v("dz", F_DZ, Real(100)); din.dpres = v("dpres", F_DPRES, Real(1000));
  din.inv_exner = v("inv_exner", F_INV_EXNER, 1);
  din.qv_prev = v("qv_prev", F_QV_PREV, Real(1e-4)); din.t_prev = v("t_prev", F_T_PREV, Real(300));

  P3F::P3DiagnosticOutputs dout;
  dout.qv2qi_depos_tend = v("qv2qi_depos_tend", F_QV2QI_DEPOS_TEND, 0);
  dout.precip_liq_surf = view_1d("precip_liq_surf", ncol);
  dout.precip_ice_surf = view_1d("precip_ice_surf", ncol);
  for (int i = 0; i < ncol; ++i) {
    dout.precip_liq_surf(i) = in(F_PRECIP_LIQ_SURF)[int64_t(i) * strides[F_PRECIP_LIQ_SURF]];
    dout.precip_ice_surf(i) = in(F_PRECIP_ICE_SURF)[int64_t(i) * strides[F_PRECIP_ICE_SURF]];
  }
  dout.diag_eff_radius_qc = v("diag_eff_radius_qc", F_DIAG_EFF_RADIUS_QC, 0);
  dout.diag_eff_radius_qi = v("diag_eff_radius_qi", F_DIAG_EFF_RADIUS_QI, 0);
  dout.diag_eff_radius_qr = v("diag_eff_radius_qr", F_DIAG_EFF_RADIUS_QR, 0);
  dout.rho_qi = v("rho_qi", F_RHO_QI, 0);
  dout.precip_liq_flux = v("precip_liq_flux", F_PRECIP_LIQ_FLUX, 0, 1);
  dout.precip_ice_flux = v("precip_ice_flux", F_PRECIP_ICE_FLUX, 0, 1);
  dout.precip_total_tend = v("precip_total_tend", F_PRECIP_TOTAL_TEND, 0);
  dout.nevapr = v("nevapr", F_NEVAPR, 0);

  P3F::P3HistoryOnly hist;
  hist.liq_ice_exchange = v("liq_ice_exchange", F_LIQ_ICE_EXCHANGE, 0);
  hist.vap_liq_exchange = v("vap_liq_exchange", F_VAP_LIQ_EXCHANGE, 0);
  hist.vap_ice_exchange = v("vap_ice_exchange", F_VAP_ICE_EXCHANGE, 0);

  P3F::P3Infrastructure infra;
  infra.dt = Real(dt); infra.it = 1; infra.its = 0; infra.ite = ncol - 1; infra.kts = 0; infra.kte = nlev - 1;
  infra.predictNc = predictNc != 0; infra.prescribedCCN = prescribedCCN != 0;
  typename P3F::view_2d<Real> col_loc("col_location", ncol, 3);
  for (int i = 0; i < ncol; ++i) { col_loc(i, 0) = Real(i + 1); col_loc(i, 1) = 0; col_loc(i, 2) = 0; }
  infra.col_location = col_loc;

  const int nk_pack = ekat::npack<Spack>(nlev);
  const auto policy = ekat::ExeSpaceUtils<typename P3F::KT::ExeSpace>::get_default_team_policy(ncol, nk_pack);
  ekat::WorkspaceManager<Spack, typename P3F::KT::Device> wsm(nk_pack, 52, policy);

  const Int us = P3F::p3_main(ro, prog, din, dout, infra, hist, h->t, wsm, ncol, nlev);

  auto back = [&](const view_2d& vw, int f, int extra = 0) { from_view(vw, out(f), strides[f], ncol, nlev + extra); };
  back(prog.qc, F_QC); back(prog.nc, F_NC); back(prog.qr, F_QR); back(prog.nr, F_NR); back(prog.qi, F_QI);
  back(prog.qm, F_QM); back(prog.ni, F_NI); back(prog.bm, F_BM); back(prog.qv, F_QV); back(prog.th, F_TH);
  back(dout.qv2qi_depos_tend, F_QV2QI_DEPOS_TEND);
  back(dout.diag_eff_radius_qc, F_DIAG_EFF_RADIUS_QC); back(dout.diag_eff_radius_qi, F_DIAG_EFF_RADIUS_QI);
  back(dout.diag_eff_radius_qr, F_DIAG_EFF_RADIUS_QR); back(dout.rho_qi, F_RHO_QI);
  back(dout.precip_liq_flux, F_PRECIP_LIQ_FLUX, 1); back(dout.precip_ice_flux, F_PRECIP_ICE_FLUX, 1);
  back(dout.precip_total_tend, F_PRECIP_TOTAL_TEND); back(dout.nevapr, F_NEVAPR);
  back(hist.liq_ice_exchange, F_LIQ_ICE_EXCHANGE); back(hist.vap_liq_exchange, F_VAP_LIQ_EXCHANGE);
  back(hist.vap_ice_exchange, F_VAP_ICE_EXCHANGE);
  for (int i = 0; i < ncol; ++i) {
    out(F_PRECIP_LIQ_SURF)[int64_t(i) * strides[F_PRECIP_LIQ_SURF]] = dout.precip_liq_surf(i);
    out(F_PRECIP_ICE_SURF)[int64_t(i) * strides[F_PRECIP_ICE_SURF]] = dout.precip_ice_surf(i);
  }
#ifdef _OPENMP
  omp_set_num_threads(saved_threads);
#endif
  return double(us) * 1e-6;
}

}  // extern "C"
