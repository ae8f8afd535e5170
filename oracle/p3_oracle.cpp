// ORACLE — TEST INFRASTRUCTURE ONLY (see p3_oracle_types.hpp header).
// extern "C" surface of the CPU restatement, loaded with ctypes by tests/, smoke() and
// bench.py's cpu_baseline leg.  Build: `make -C oracle` -> oracle/libp3oracle.so
#include "p3_oracle_main.hpp"

#include <chrono>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <string>
#ifdef _OPENMP
#include <omp.h>
#endif

using namespace p3o;

double p3o::pm::g_scale[p3o::pm::NUM_FN] = {1, 1, 1, 1, 1, 1, 1, 1};
int p3o::pm::g_shared = 0;

namespace {

// Field order shared by the oracle and the product C ABI (include/p3_b200.h, P3B200_F_*).
enum Field {
  F_QC, F_NC, F_QR, F_NR, F_QI, F_QM, F_NI, F_BM, F_QV, F_TH,
  F_NC_NUCEAT_TEND, F_NCCN, F_NI_ACTIVATED, F_INV_QC_RELVAR, F_CLD_FRAC_I, F_CLD_FRAC_L, F_CLD_FRAC_R,
  F_PRES, F_DZ, F_DPRES, F_INV_EXNER, F_QV_PREV, F_T_PREV,
  F_QV2QI_DEPOS_TEND, F_DIAG_EFF_RADIUS_QC, F_DIAG_EFF_RADIUS_QI, F_DIAG_EFF_RADIUS_QR, F_RHO_QI,
  F_PRECIP_LIQ_FLUX, F_PRECIP_ICE_FLUX, F_PRECIP_TOTAL_TEND, F_NEVAPR,
  F_LIQ_ICE_EXCHANGE, F_VAP_LIQ_EXCHANGE, F_VAP_ICE_EXCHANGE,
  F_PRECIP_LIQ_SURF, F_PRECIP_ICE_SURF, F_NUM
};

template <typename Real>
struct TablesHolder { P3LookupTables<Real> t; };

struct OracleTables {
  TablesHolder<double>* d = nullptr;
  TablesHolder<float>*  f = nullptr;
};

template <typename Real>
P3Runtime<Real> make_runtime(const double* r) {
  P3Runtime<Real> ro;
  if (!r) return ro;
  ro.max_total_ni = r[0]; ro.autoconversion_prefactor = r[1]; ro.autoconversion_qc_exponent = r[2];
  ro.autoconversion_nc_exponent = r[3]; ro.autoconversion_radius = r[4]; ro.accretion_prefactor = r[5];
  ro.accretion_qc_exponent = r[6]; ro.accretion_qr_exponent = r[7]; ro.rain_selfcollection_prefactor = r[8];
  ro.rain_selfcollection_breakup_diameter = r[9]; ro.constant_mu_rain = r[10]; ro.spa_ccn_to_nc_factor = r[11];
  ro.cldliq_to_ice_collection_factor = r[12]; ro.rain_to_ice_collection_factor = r[13];
  ro.min_rime_rho = r[14]; ro.max_rime_rho = r[15]; ro.immersion_freezing_exponent = r[16];
  ro.deposition_nucleation_exponent = r[17]; ro.ice_sedimentation_factor = r[18];
  ro.do_ice_production = r[19] != 0.0;
  return ro;
}

template <typename Real>
double run_main(const P3LookupTables<Real>& tb, void* const* fields, const int64_t* strides,
                int ncol, int nlev, double dt, int predictNc, int prescribedCCN,
                const double* runtime, int* diag, int nthreads) {
  const P3Runtime<Real> ro = make_runtime<Real>(runtime);
  auto col_ptr = [&](int f, int i) { return static_cast<Real*>(fields[f]) + int64_t(i)*strides[f]; };
  const auto start = std::chrono::steady_clock::now();
#ifdef _OPENMP
  if (nthreads <= 0) nthreads = omp_get_max_threads();
#pragma omp parallel num_threads(nthreads)
#endif
  {
    Workspace<Real> w(nlev);   // one workspace per thread = one team per column, 1 thread per team
#ifdef _OPENMP
#pragma omp for schedule(dynamic, 16)
#endif
    for (int i = 0; i < ncol; ++i) {
      Column<Real> c;
      c.qc = col_ptr(F_QC,i); c.nc = col_ptr(F_NC,i); c.qr = col_ptr(F_QR,i); c.nr = col_ptr(F_NR,i);
      c.qi = col_ptr(F_QI,i); c.qm = col_ptr(F_QM,i); c.ni = col_ptr(F_NI,i); c.bm = col_ptr(F_BM,i);
      c.qv = col_ptr(F_QV,i); c.th = col_ptr(F_TH,i);
      c.nc_nuceat_tend = col_ptr(F_NC_NUCEAT_TEND,i); c.nccn = col_ptr(F_NCCN,i);
      c.ni_activated = col_ptr(F_NI_ACTIVATED,i); c.inv_qc_relvar = col_ptr(F_INV_QC_RELVAR,i);
      c.cld_frac_i = col_ptr(F_CLD_FRAC_I,i); c.cld_frac_l = col_ptr(F_CLD_FRAC_L,i); c.cld_frac_r = col_ptr(F_CLD_FRAC_R,i);
      c.pres = col_ptr(F_PRES,i); c.dz = col_ptr(F_DZ,i); c.dpres = col_ptr(F_DPRES,i);
      c.inv_exner = col_ptr(F_INV_EXNER,i); c.qv_prev = col_ptr(F_QV_PREV,i); c.t_prev = col_ptr(F_T_PREV,i);
      c.qv2qi_depos_tend = col_ptr(F_QV2QI_DEPOS_TEND,i);
      c.diag_eff_radius_qc = col_ptr(F_DIAG_EFF_RADIUS_QC,i); c.diag_eff_radius_qi = col_ptr(F_DIAG_EFF_RADIUS_QI,i);
      c.diag_eff_radius_qr = col_ptr(F_DIAG_EFF_RADIUS_QR,i); c.rho_qi = col_ptr(F_RHO_QI,i);
      c.precip_liq_flux = col_ptr(F_PRECIP_LIQ_FLUX,i); c.precip_ice_flux = col_ptr(F_PRECIP_ICE_FLUX,i);
      c.precip_total_tend = col_ptr(F_PRECIP_TOTAL_TEND,i); c.nevapr = col_ptr(F_NEVAPR,i);
      c.liq_ice_exchange = col_ptr(F_LIQ_ICE_EXCHANGE,i); c.vap_liq_exchange = col_ptr(F_VAP_LIQ_EXCHANGE,i);
      c.vap_ice_exchange = col_ptr(F_VAP_ICE_EXCHANGE,i);
      c.precip_liq_surf = col_ptr(F_PRECIP_LIQ_SURF,i); c.precip_ice_surf = col_ptr(F_PRECIP_ICE_SURF,i);
      // the workspace is reused across columns exactly like the reference's WorkspaceManager slots;
      // p3_main_init re-initialises every array that is read before being written.
      Main<Real>::p3_main_column(ro, c, static_cast<Real>(dt), predictNc != 0, prescribedCCN != 0, tb, nlev, w,
                                 diag ? diag + size_t(i)*D_NUM : nullptr);
    }
  }
  const auto finish = std::chrono::steady_clock::now();
  return std::chrono::duration<double>(finish - start).count();
}

} // namespace

extern "C" {

// p3_table_ice_impl.hpp:18-78 — parse the text table.  ice: [5][4][50][12]; coll_raw: [5][4][50][30][2]
// holds the RAW (not log10) collection values so callers can apply log10 themselves.
// Returns 0, or -1 (cannot open), -2 (bad header), -3 (bad version), -4 (short file).
int p3o_read_ice_table_text(const char* path, const char* expect_version, double* ice, double* coll_raw) {
  std::ifstream in(path);
  if (!in) return -1;
  std::string version, version_val;
  in >> version >> version_val;
  if (version != "VERSION") return -2;
  if (expect_version && version_val != expect_version) return -3;
  double dum_s; int dum_i;
  for (int jj = 0; jj < 5; ++jj) for (int ii = 0; ii < 4; ++ii) {
    for (int i = 0; i < 50; ++i) {
      in >> dum_i >> dum_i;
      int j_idx = 0;
      for (int j = 0; j < 15; ++j) {
        in >> dum_s;
        if (j > 1 && j != 10) ice[((jj*4+ii)*50+i)*12 + j_idx++] = dum_s;
      }
    }
    for (int i = 0; i < 50; ++i) for (int j = 0; j < 30; ++j) {
      in >> dum_i >> dum_i;
      int k_idx = 0;
      for (int k = 0; k < 6; ++k) {
        in >> dum_s;
        if (k == 3 || k == 4) coll_raw[(((jj*4+ii)*50+i)*30+j)*2 + k_idx++] = dum_s;
      }
    }
  }
  return in ? 0 : -4;
}

// components/eam/src/physics/p3/scream/micro_p3.F90:218-341 (p3_init_b): rain fall-speed and
// ventilation tables by numerical integration over the PSD.  Outputs are [300][10] row-major
// (C++ order after the transpose in p3_table3_impl.hpp:111-119) and mu_r[150].
void p3o_gen_rain_tables(double* vn, double* vm, double* revap, double* mu_r_table) {
  const double pi = Constants<double>::Pi;
  const double thrd = 1.0/3.0, sxth = 1.0/6.0, piov6 = pi*sxth;
  const double mu_r_constant = 1.0;
  for (int i = 0; i < 150; ++i) mu_r_table[i] = mu_r_constant;
  for (int ii = 1; ii <= 10; ++ii) {
    const double mu_r = mu_r_constant;
    for (int jj = 1; jj <= 300; ++jj) {
      double dm;
      if (jj <= 20) dm = (double(jj)*10. - 5.)*1.e-6;
      else          dm = (double(jj-20)*30. + 195.)*1.e-6;
      const double lamr = (mu_r+1.)/dm;
      double dum1 = 0, dum2 = 0, dum3 = 0, dum4 = 0, dum5 = 0;
      const double dd = 2.;
      for (int kk = 1; kk <= 10000; ++kk) {
        const double dia = (double(kk)*dd - dd*0.5)*1.e-6;
        double amg = piov6*997.*(dia*dia*dia);
        amg = amg*1000.;
        double vt;
        if (dia*1.e+6 <= 134.43)       vt = 4.5795e+3*std::pow(amg, 2.*thrd);
        else if (dia*1.e+6 < 1511.64)  vt = 4.962e+1*std::pow(amg, thrd);
        else if (dia*1.e+6 < 3477.84)  vt = 1.732e+1*std::pow(amg, sxth);
        else                           vt = 9.17;
        // left-to-right like the Fortran expression: ((vt*10**x)*exp(..))*dd*1e-6  (F90:321-325)
        const double ex = std::exp(-lamr*dia);
        dum1 = dum1 + vt*std::pow(10., mu_r*std::log10(dia)+4.*mu_r)*ex*dd*1.e-6;
        dum2 = dum2 + std::pow(10., mu_r*std::log10(dia)+4.*mu_r)*ex*dd*1.e-6;
        dum3 = dum3 + vt*std::pow(10., (mu_r+3.)*std::log10(dia)+4.*mu_r)*ex*dd*1.e-6;
        dum4 = dum4 + std::pow(10., (mu_r+3.)*std::log10(dia)+4.*mu_r)*ex*dd*1.e-6;
        dum5 = dum5 + std::pow(vt*dia, 0.5)*std::pow(10., (mu_r+1.)*std::log10(dia)+3.*mu_r)*ex*dd*1.e-6;
      }
      dum2 = std::max(dum2, 1.e-30);
      dum4 = std::max(dum4, 1.e-30);
      dum5 = std::max(dum5, 1.e-30);
      vn[(jj-1)*10 + (ii-1)]    = dum1/dum2;
      vm[(jj-1)*10 + (ii-1)]    = dum3/dum4;
      revap[(jj-1)*10 + (ii-1)] = std::pow(10., std::log10(dum5)+(mu_r+1.)*std::log10(lamr)-(3.*mu_r));
    }
    if (ii == 1) {   // mu_r is constant, so the other 9 columns repeat column 1 bit-for-bit
      for (int jj = 0; jj < 300; ++jj) for (int c2 = 1; c2 < 10; ++c2) {
        vn[jj*10+c2] = vn[jj*10]; vm[jj*10+c2] = vm[jj*10]; revap[jj*10+c2] = revap[jj*10];
      }
      break;
    }
  }
}

// Build both precisions of P3LookupTables from host arrays (ice[12000], coll_raw[60000],
// vn/vm/revap[3000], mu_r[150]).  dnu literals: p3_table3_impl.hpp:120-135.
void* p3o_tables_create(const double* ice, const double* coll_raw, const double* vn, const double* vm,
                        const double* revap, const double* mu_r) {
  auto* o = new OracleTables;
  o->d = new TablesHolder<double>;
  o->f = new TablesHolder<float>;
  static const double dnu[16] = {0.000,-0.557,-0.430,-0.307,-0.186,-0.067,-0.050,-0.167,-0.282,-0.397,
                                 -0.512,-0.626,-0.739,-0.853,-0.966,-0.966};
  auto fill = [&](auto& t) {
    using R = std::remove_reference_t<decltype(t.mu_r_table_vals[0])>;
    for (int i = 0; i < 150; ++i) t.mu_r_table_vals[i] = R(mu_r[i]);
    for (int i = 0; i < 3000; ++i) {
      (&t.vn_table_vals[0][0])[i] = R(vn[i]); (&t.vm_table_vals[0][0])[i] = R(vm[i]);
      (&t.revap_table_vals[0][0])[i] = R(revap[i]);
    }
    for (int i = 0; i < 12000; ++i) (&t.ice_table_vals[0][0][0][0])[i] = R(ice[i]);
    for (int i = 0; i < 60000; ++i) (&t.collect_table_vals[0][0][0][0][0])[i] = R(std::log10(coll_raw[i]));
    for (int i = 0; i < 16; ++i) t.dnu_table_vals[i] = R(dnu[i]);
  };
  fill(o->d->t); fill(o->f->t);
  return o;
}

void p3o_tables_destroy(void* h) {
  auto* o = static_cast<OracleTables*>(h);
  if (!o) return;
  delete o->d; delete o->f; delete o;
}

int p3o_num_fields() { return F_NUM; }
int p3o_num_diag()   { return D_NUM; }
int p3o_max_threads() {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

// p3_main over ncol columns (host pointers, level-contiguous, per-field column stride in elements).
// real_bytes = 8 (double) or 4 (float).  runtime = 20 doubles (P3Runtime order) or NULL for defaults.
// diag = ncol*D_NUM ints or NULL.  Returns elapsed seconds around the column loop
// (mirrors p3_main_impl.hpp:105,326-330), or a negative error code.
double p3o_main(void* tables, int real_bytes, void* const* fields, const int64_t* strides,
                int ncol, int nlev, double dt, int predictNc, int prescribedCCN,
                const double* runtime, int* diag, int nthreads) {
  auto* o = static_cast<OracleTables*>(tables);
  if (!o) return -1.0;
  if (real_bytes == 8) return run_main<double>(o->d->t, fields, strides, ncol, nlev, dt, predictNc, prescribedCCN, runtime, diag, nthreads);
  if (real_bytes == 4) return run_main<float>(o->f->t, fields, strides, ncol, nlev, dt, predictNc, prescribedCCN, runtime, diag, nthreads);
  return -2.0;
}

} // extern "C"

// Conditioning probe: relative scale applied to the result of pow, exp, log, log10, cbrt, tgamma,
// tanh, expm1 (in this order).  nullptr restores the exact libm values.  Not thread safe: call between runs.
extern "C" void p3o_set_math_scale(const double* scale8) {
  for (int i = 0; i < p3o::pm::NUM_FN; ++i) p3o::pm::g_scale[i] = scale8 ? scale8[i] : 1.0;
}

// 0: host libm (default), 1: the deterministic transcendentals shared with the kernels (p3_math.hpp)
extern "C" void p3o_set_math_mode(int shared) { p3o::pm::g_shared = shared ? 1 : 0; }
extern "C" int p3o_get_math_mode() { return p3o::pm::g_shared; }

// Evaluate one transcendental family (pm::Fn order) of the shared header on arrays: tests/test_cpu_shared_math.py
// measures its error against mpmath.  real_bytes 4 evaluates the binary32 overloads (inputs/outputs still double).
extern "C" void p3o_math_eval(int fn, int real_bytes, int n, const double* x, const double* y, double* out) {
  for (int i = 0; i < n; ++i) {
    if (real_bytes == 8) {
      const double a = x[i], b = y ? y[i] : 0.0;
      switch (fn) {
        case p3o::pm::POW: out[i] = p3math::pow(a, b); break;
        case p3o::pm::EXP: out[i] = p3math::exp(a); break;
        case p3o::pm::LOG: out[i] = p3math::log(a); break;
        case p3o::pm::LOG10: out[i] = p3math::log10(a); break;
        case p3o::pm::CBRT: out[i] = p3math::cbrt(a); break;
        case p3o::pm::TGAMMA: out[i] = p3math::tgamma(a); break;
        case p3o::pm::TANH: out[i] = p3math::tanh(a); break;
        default: out[i] = p3math::expm1(a); break;
      }
    } else {
      const float a = (float)x[i], b = y ? (float)y[i] : 0.0f;
      switch (fn) {
        case p3o::pm::POW: out[i] = p3math::pow(a, b); break;
        case p3o::pm::EXP: out[i] = p3math::exp(a); break;
        case p3o::pm::LOG: out[i] = p3math::log(a); break;
        case p3o::pm::LOG10: out[i] = p3math::log10(a); break;
        case p3o::pm::CBRT: out[i] = p3math::cbrt(a); break;
        case p3o::pm::TGAMMA: out[i] = p3math::tgamma(a); break;
        case p3o::pm::TANH: out[i] = p3math::tanh(a); break;
        default: out[i] = p3math::expm1(a); break;
      }
    }
  }
}
