// ORACLE — TEST INFRASTRUCTURE ONLY (see p3_oracle_types.hpp header).
// Column driver of P3: p3_main_init / part1 / part2 / sedimentation / homogeneous freezing / part3,
// restated for one column with scalar semantics and the same statement order as the reference.
// Citations: p3/impl/<file>:<lines> in /root/reference/components/eamxx/src/physics.
#ifndef P3_ORACLE_MAIN_HPP
#define P3_ORACLE_MAIN_HPP

#include "p3_oracle_procs.hpp"
#include <vector>

namespace p3o {

// The 44 per-column temporaries of p3_main_internal (p3_main_impl.hpp:120-163).
template <typename Real>
struct Workspace {
  int nk;
  std::vector<Real> store;
  Real *mu_r, *T_atm, *lamr, *logn0r, *nu, *cdist, *cdist1, *cdistr,
       *inv_cld_frac_i, *inv_cld_frac_l, *inv_cld_frac_r, *qc_incld, *qr_incld, *qi_incld, *qm_incld,
       *nc_incld, *nr_incld, *ni_incld, *bm_incld,
       *inv_dz, *inv_rho, *ze_ice, *ze_rain, *prec, *rho,
       *rhofacr, *rhofaci, *acn, *qv_sat_l, *qv_sat_i, *sup, *qv_supersat_i,
       *tmparr1, *exner, *diag_equiv_reflectivity, *diag_vm_qi, *diag_diam_qi,
       *pratot, *prctot, *qtend_ignore, *ntend_ignore, *mu_c, *lamc, *qr_evap_tend;
  // sedimentation scratch (p3_cloud_sed_impl.hpp:42-46, p3_ice_sed_impl.hpp:98-102)
  Real *V_q, *V_n, *flux[4];
  explicit Workspace(int nk_) : nk(nk_), store(size_t(50)*nk_, Real(0)) {
    Real* p = store.data();
    Real** all[] = {&mu_r,&T_atm,&lamr,&logn0r,&nu,&cdist,&cdist1,&cdistr,&inv_cld_frac_i,&inv_cld_frac_l,
      &inv_cld_frac_r,&qc_incld,&qr_incld,&qi_incld,&qm_incld,&nc_incld,&nr_incld,&ni_incld,&bm_incld,
      &inv_dz,&inv_rho,&ze_ice,&ze_rain,&prec,&rho,&rhofacr,&rhofaci,&acn,&qv_sat_l,&qv_sat_i,&sup,
      &qv_supersat_i,&tmparr1,&exner,&diag_equiv_reflectivity,&diag_vm_qi,&diag_diam_qi,&pratot,&prctot,
      &qtend_ignore,&ntend_ignore,&mu_c,&lamc,&qr_evap_tend,&V_q,&V_n,&flux[0],&flux[1],&flux[2],&flux[3]};
    for (auto a : all) { *a = p; p += nk; }
  }
};

template <typename Real>
struct Main {
  using P = Procs<Real>;
  using C = Constants<Real>;
  using Runtime = P3Runtime<Real>;
  using Tables  = P3LookupTables<Real>;
  using Col = Column<Real>;
  using WS  = Workspace<Real>;
  using Tend = typename P::Tend;
  static Real sp(double x) { return static_cast<Real>(x); }

  // p3_main_impl.hpp:21-72
  static void p3_main_init(int nk, const Col& c, WS& w) {
    *c.precip_liq_surf = 0;
    *c.precip_ice_surf = 0;
    for (int k = 0; k < nk; ++k) {
      w.diag_equiv_reflectivity[k] = -99;
      w.ze_ice[k]  = sp(1.e-22);
      w.ze_rain[k] = sp(1.e-22);
      c.diag_eff_radius_qc[k] = sp(10.e-6);
      c.diag_eff_radius_qi[k] = sp(25.e-6);
      c.diag_eff_radius_qr[k] = sp(500.e-6);
      w.inv_cld_frac_i[k] = 1 / c.cld_frac_i[k];
      w.inv_cld_frac_l[k] = 1 / c.cld_frac_l[k];
      w.inv_cld_frac_r[k] = 1 / c.cld_frac_r[k];
      w.exner[k] = 1 / c.inv_exner[k];
      w.T_atm[k] = c.th[k] * w.exner[k];
      c.qv[k]    = P::max(c.qv[k], 0);
      w.inv_dz[k] = 1 / c.dz[k];
      // zero_init list, p3_main_impl.hpp:211-217
      Real* z[] = {w.mu_r, w.lamr, w.logn0r, w.nu, w.cdist, w.cdist1, w.cdistr,
        w.qc_incld, w.qr_incld, w.qi_incld, w.qm_incld, w.nc_incld, w.nr_incld, w.ni_incld, w.bm_incld,
        w.inv_rho, w.prec, w.rho, w.rhofacr, w.rhofaci, w.acn, w.qv_sat_l, w.qv_sat_i, w.sup, w.qv_supersat_i,
        w.tmparr1, w.qtend_ignore, w.ntend_ignore, w.mu_c, w.lamc,
        c.rho_qi, c.qv2qi_depos_tend, c.precip_total_tend, c.nevapr, c.precip_liq_flux, c.precip_ice_flux};
      for (Real* a : z) a[k] = 0;
    }
    // NOTE: with pack size 1 the reference leaves precip_*_flux[nk] untouched
    // here; its own test bridge pre-zeroes it (p3_functions_f90.cpp:1941-1945).  The oracle follows
    // the reference: element [nk] is NOT zeroed by p3_main.
  }

  // p3_main_impl_part1.hpp:21-185
  static void p3_main_part1(int nk, bool predictNc, bool do_prescribed_CCN, Real dt, const Col& c, WS& w,
                            bool& nucleationPossible, bool& hydrometeorsPresent, const Runtime& ro, int* err) {
    const Real g = C::gravit, rho_1000mb = C::RHO_1000MB, rho_600mb = C::RHO_600MB, rho_h2o = C::RHO_H2O;
    const Real nccnst = C::NCCNST, T_zerodegc = C::T_zerodegc, qsmall = C::QSMALL, inv_cp = C::INV_CP;
    const Real latvap = C::LatVap, latice = C::LatIce;
    const Real spa_ccn_to_nc_factor = ro.spa_ccn_to_nc_factor;
    nucleationPossible = false;
    hydrometeorsPresent = false;
    for (int k = 0; k < nk; ++k) {
      w.rho[k]      = c.dpres[k]/c.dz[k] / g;
      w.inv_rho[k]  = 1 / w.rho[k];
      w.qv_sat_l[k] = P::qv_sat_dry(w.T_atm[k], c.pres[k], false, err);
      w.qv_sat_i[k] = P::qv_sat_dry(w.T_atm[k], c.pres[k], true,  err);
      w.qv_supersat_i[k] = c.qv[k] / w.qv_sat_i[k] - 1;
      w.rhofacr[k] = pm::pow(rho_1000mb * w.inv_rho[k], sp(.54));
      w.rhofaci[k] = pm::pow(rho_600mb * w.inv_rho[k], sp(.54));
      const Real dum = sp(1.496e-6) * pm::pow(w.T_atm[k], sp(1.5)) / (w.T_atm[k] + 120);
      w.acn[k] = g * rho_h2o / (18 * dum);
      if (w.T_atm[k] < T_zerodegc && w.qv_supersat_i[k] >= sp(-0.05)) nucleationPossible = true;

      bool drymass = c.qc[k] < qsmall;
      if (drymass) {
        c.qv[k] = c.qv[k] + c.qc[k];
        c.th[k] = c.th[k] - c.inv_exner[k] * c.qc[k] * latvap * inv_cp;
        c.qc[k] = 0; c.nc[k] = 0;
      } else {
        hydrometeorsPresent = true;
        if (do_prescribed_CCN)
          c.nc[k] = P::max(c.nc[k], spa_ccn_to_nc_factor * c.nccn[k] / w.inv_cld_frac_l[k]);
        else if (predictNc)
          c.nc[k] = P::max(c.nc[k] + c.nc_nuceat_tend[k] * dt, sp(0.0));
        else
          c.nc[k] = nccnst * w.inv_rho[k];
      }

      drymass = c.qr[k] < qsmall;
      if (drymass) {
        c.qv[k] = c.qv[k] + c.qr[k];
        c.th[k] = c.th[k] - c.inv_exner[k] * c.qr[k] * latvap * inv_cp;
        c.qr[k] = 0; c.nr[k] = 0;
      } else {
        hydrometeorsPresent = true;
      }

      drymass = (c.qi[k] < qsmall || (c.qi[k] < sp(1.e-8) && w.qv_supersat_i[k] < sp(-0.1)));
      if (drymass) {
        c.qv[k] = c.qv[k] + c.qi[k];
        c.th[k] = c.th[k] - c.inv_exner[k] * c.qi[k] * (latvap+latice) * inv_cp;
        c.qi[k] = 0; c.ni[k] = 0; c.qm[k] = 0; c.bm[k] = 0;
      } else {
        hydrometeorsPresent = true;
      }

      drymass = (c.qi[k] >= qsmall && c.qi[k] < sp(1.e-8) && w.T_atm[k] >= T_zerodegc);
      if (drymass) {
        c.qr[k] = c.qr[k] + c.qi[k];
        c.th[k] = c.th[k] - c.inv_exner[k] * c.qi[k] * latice * inv_cp;
        c.qi[k] = 0; c.ni[k] = 0; c.qm[k] = 0; c.bm[k] = 0;
      }

      w.T_atm[k] = c.th[k] * w.exner[k];

      P::calculate_incloud_mixingratios(
        c.qc[k], c.qr[k], c.qi[k], c.qm[k], c.nc[k], c.nr[k], c.ni[k], c.bm[k],
        w.inv_cld_frac_l[k], w.inv_cld_frac_i[k], w.inv_cld_frac_r[k],
        w.qc_incld[k], w.qr_incld[k], w.qi_incld[k], w.qm_incld[k], w.nc_incld[k], w.nr_incld[k],
        w.ni_incld[k], w.bm_incld[k]);
    }
  }

  // p3_find_impl.hpp:57-92 / 17-52 (team_size()==1 branch; kdir = -1 in p3_main)
  static int find_top(const Real* v, Real small, int kbot, int ktop, int kdir, bool& log_present) {
    log_present = false;
    int k_xtop = 0;
    for (int k = ktop; k != kbot - kdir; k -= kdir) {
      if (v[k] < small) continue;
      k_xtop = k; log_present = true; break;
    }
    return k_xtop;
  }
  static int find_bottom(const Real* v, Real small, int kbot, int ktop, int kdir, bool& log_present) {
    log_present = false;
    int k_xbot = 0;
    for (int k = kbot; k != ktop + kdir; k += kdir) {
      if (v[k] < small) continue;
      k_xbot = k; log_present = true; break;
    }
    return k_xbot;
  }

  // p3_upwind_impl.hpp:18-91 (pack size 1: k_top_pack == k_top, the "single" block handles the
  // inflow cell with zero flux from outside the range, then the remaining cells use flux(k+kdir)).
  template <int nfield>
  static void calc_first_order_upwind_step(const Real* rho, const Real* inv_rho, const Real* inv_dz,
      int /*nk*/, int k_bot, int k_top, int kdir, Real dt_sub, Real* const* flux, Real* const* V, Real* const* r) {
    int kmin = (kdir == 1 ? k_bot : k_top);
    int kmax = (kdir == 1 ? k_top : k_bot) + 1;     // [kmin, kmax)
    for (int k = kmin; k < kmax; ++k)
      for (int f = 0; f < nfield; ++f)
        flux[f][k] = V[f][k] * r[f][k] * rho[k];
    {
      const int k = k_top;
      for (int f = 0; f < nfield; ++f) {
        const Real flux_pkdir = 0;
        const Real fluxdiv = (flux_pkdir - flux[f][k]) * inv_dz[k];
        r[f][k] += fluxdiv * dt_sub * inv_rho[k];
      }
    }
    if (kdir == 1) --kmax; else ++kmin;
    for (int k = kmin; k < kmax; ++k)
      for (int f = 0; f < nfield; ++f) {
        const Real flux_pkdir = flux[f][k+kdir];
        const Real fluxdiv = (flux_pkdir - flux[f][k]) * inv_dz[k];
        r[f][k] += fluxdiv * dt_sub * inv_rho[k];
      }
  }

  // p3_upwind_impl.hpp:97-129
  template <int nfield>
  static void generalized_sedimentation(const Real* rho, const Real* inv_rho, const Real* inv_dz,
      int nk, int k_qxtop, int& k_qxbot, int kbot, int kdir, Real Co_max, Real& dt_left, Real& prt_accum,
      Real* const* fluxes, Real* const* Vs, Real* const* rs, int* err) {
    if (!(Co_max >= 0) && err) *err |= 4;
    const int tmpint1 = static_cast<int>(Co_max + 1);
    const Real dt_sub = dt_left/tmpint1;
    const int k_temp = (k_qxbot == kbot) ? k_qxbot : k_qxbot - kdir;
    calc_first_order_upwind_step<nfield>(rho, inv_rho, inv_dz, nk, k_temp, k_qxtop, kdir, dt_sub, fluxes, Vs, rs);
    if (k_qxbot == kbot) prt_accum += fluxes[0][kbot] * dt_sub;
    else k_qxbot -= kdir;
    dt_left -= dt_sub;
  }

  // p3_main_impl_part2.hpp:21-526
  static void p3_main_part2(int nk, bool predictNc, bool do_prescribed_CCN, Real dt, Real inv_dt,
      const Tables& tb, const Col& c, WS& w, bool& hydrometeorsPresent, const Runtime& ro, int* err) {
    const Real qsmall = C::QSMALL, nsmall = C::NSMALL, T_zerodegc = C::T_zerodegc;
    const Real f1r = C::f1r, f2r = C::f2r, nmltratio = C::nmltratio, inv_cp = C::INV_CP;
    const Real latvap = C::LatVap, latice = C::LatIce;
    const Real max_total_ni = ro.max_total_ni;
    const bool do_ice_production = ro.do_ice_production;
    hydrometeorsPresent = false;

    for (int k = 0; k < nk; ++k) {
      const bool skip_all = (c.qc[k] < qsmall && c.qr[k] < qsmall && c.qi[k] < qsmall &&
                             w.T_atm[k] < T_zerodegc && w.qv_supersat_i[k] < sp(-0.05));
      if (skip_all) continue;

      Tend t;
      Real mu=0, dv=0, sc=0, dqsdt=0, dqsidt=0, ab=0, abi=0, kap=0, eii=0;
      Real table_val_qi_fallspd=0, table_val_ni_self_collect=0, table_val_qc2qi_collect=0,
           table_val_qi2qr_melting=0, table_val_nr_collect=0, table_val_qr2qi_collect=0,
           table_val_ni_lammax=0, table_val_ni_lammin=0, table_val_qi2qr_vent_melt=0;
      Real vtrmi1=0, rho_qm_cloud=400, epsi=0, epsr=0, epsc=0, epsi_tot=0;
      bool wetgrowth = false;

      const bool skip_micro = !(w.qc_incld[k] >= qsmall || w.qr_incld[k] >= qsmall || w.qi_incld[k] >= qsmall);
      if (!skip_micro) {
        P::get_time_space_phys_variables(w.T_atm[k], c.pres[k], w.rho[k], w.qv_sat_l[k], w.qv_sat_i[k],
                                         mu, dv, sc, dqsdt, dqsidt, ab, abi, kap, eii);
        P::get_cloud_dsd2(w.qc_incld[k], w.nc_incld[k], w.mu_c[k], w.rho[k], w.nu[k], w.lamc[k], w.cdist[k], w.cdist1[k]);
        c.nc[k] = w.nc_incld[k] * c.cld_frac_l[k];
        P::get_rain_dsd2(w.qr_incld[k], w.nr_incld[k], w.mu_r[k], w.lamr[k], ro);
        P::get_cdistr_logn0r(w.qr_incld[k], w.nr_incld[k], w.mu_r[k], w.lamr[k], w.cdistr[k], w.logn0r[k]);
        c.nr[k] = w.nr_incld[k] * c.cld_frac_r[k];
        P::impose_max_total_ni(w.ni_incld[k], max_total_ni, w.inv_rho[k]);

        const bool qi_gt_small = w.qi_incld[k] >= qsmall;
        if (qi_gt_small) {
          w.ni_incld[k] = P::max(w.ni_incld[k], nsmall);
          w.nr_incld[k] = P::max(w.nr_incld[k], nsmall);
          const Real rhop = P::calc_bulk_rho_rime(w.qi_incld[k], w.qm_incld[k], w.bm_incld[k], ro);
          c.qm[k] = w.qm_incld[k]*c.cld_frac_i[k];
          c.bm[k] = w.bm_incld[k]*c.cld_frac_i[k];
          TableIce<Real> table_ice;
          P::lookup_ice(w.qi_incld[k], w.ni_incld[k], w.qm_incld[k], rhop, table_ice);
          TableRain<Real> table_rain;
          P::lookup_rain(w.qr_incld[k], w.nr_incld[k], table_rain);
          table_val_qi_fallspd      = P::apply_table_ice(1, tb, table_ice);
          table_val_ni_self_collect = P::apply_table_ice(2, tb, table_ice);
          table_val_qc2qi_collect   = P::apply_table_ice(3, tb, table_ice);
          table_val_qi2qr_melting   = P::apply_table_ice(4, tb, table_ice);
          table_val_ni_lammax       = P::apply_table_ice(6, tb, table_ice);
          table_val_ni_lammin       = P::apply_table_ice(7, tb, table_ice);
          table_val_qi2qr_vent_melt = P::apply_table_ice(9, tb, table_ice);
          if (w.qr_incld[k] >= qsmall) {
            table_val_nr_collect    = P::apply_table_coll(0, tb, table_ice, table_rain);
            table_val_qr2qi_collect = P::apply_table_coll(1, tb, table_ice, table_rain);
          }
          w.ni_incld[k] = P::min(w.ni_incld[k], table_val_ni_lammax*w.ni_incld[k]);
          w.ni_incld[k] = P::max(w.ni_incld[k], table_val_ni_lammin*w.ni_incld[k]);
        }

        if (do_ice_production) {
          P::ice_cldliq_collection(w.rho[k], w.T_atm[k], w.rhofaci[k], table_val_qc2qi_collect,
              w.qi_incld[k], w.qc_incld[k], w.ni_incld[k], w.nc_incld[k],
              t.qc2qi_collect_tend, t.nc_collect_tend, t.qc2qr_ice_shed_tend, t.ncshdc, ro);
          P::ice_rain_collection(w.rho[k], w.T_atm[k], w.rhofaci[k], w.logn0r[k],
              table_val_nr_collect, table_val_qr2qi_collect, w.qi_incld[k], w.ni_incld[k], w.qr_incld[k],
              t.qr2qi_collect_tend, t.nr_collect_tend, ro);
          P::ice_self_collection(w.rho[k], w.rhofaci[k], table_val_ni_self_collect, eii,
              w.qm_incld[k], w.qi_incld[k], w.ni_incld[k], t.ni_selfcollect_tend);
        }
        P::ice_melting(w.rho[k], w.T_atm[k], c.pres[k], w.rhofaci[k], table_val_qi2qr_melting,
            table_val_qi2qr_vent_melt, dv, sc, mu, kap, c.qv[k], w.qi_incld[k], w.ni_incld[k],
            t.qi2qr_melt_tend, t.ni2nr_melt_tend, err);
        if (do_ice_production) {
          P::ice_cldliq_wet_growth(w.rho[k], w.T_atm[k], c.pres[k], w.rhofaci[k], table_val_qi2qr_melting,
              table_val_qi2qr_vent_melt, dv, kap, mu, sc, c.qv[k], w.qc_incld[k], w.qi_incld[k],
              w.ni_incld[k], w.qr_incld[k], wetgrowth, t.qr2qi_collect_tend, t.qc2qi_collect_tend,
              t.qc_growth_rate, t.nr_ice_shed_tend, t.qc2qr_ice_shed_tend, err);
        }
        P::ice_relaxation_timescale(w.rho[k], w.T_atm[k], w.rhofaci[k], table_val_qi2qr_melting,
            table_val_qi2qr_vent_melt, dv, mu, sc, w.qi_incld[k], w.ni_incld[k], epsi, epsi_tot);
        P::calc_rime_density(w.T_atm[k], w.rhofaci[k], table_val_qi_fallspd, w.acn[k], w.lamc[k], w.mu_c[k],
            w.qc_incld[k], t.qc2qi_collect_tend, vtrmi1, rho_qm_cloud);
        if (do_ice_production) {
          P::cldliq_immersion_freezing(w.T_atm[k], w.lamc[k], w.mu_c[k], w.cdist1[k], w.qc_incld[k],
              c.inv_qc_relvar[k], t.qc2qi_hetero_freeze_tend, t.nc2ni_immers_freeze_tend, ro);
          P::rain_immersion_freezing(w.T_atm[k], w.lamr[k], w.mu_r[k], w.cdistr[k], w.qr_incld[k],
              t.qr2qi_immers_freeze_tend, t.nr2ni_immers_freeze_tend, ro);
        }
        P::calc_liq_relaxation_timescale(tb, w.rho[k], f1r, f2r, dv, mu, sc, w.mu_r[k], w.lamr[k],
            w.cdistr[k], w.cdist[k], w.qr_incld[k], w.qc_incld[k], epsr, epsc);
        P::evaporate_rain(w.qr_incld[k], w.qc_incld[k], w.nr_incld[k], w.qi_incld[k],
            c.cld_frac_l[k], c.cld_frac_r[k], c.qv[k], c.qv_prev[k], w.qv_sat_l[k], w.qv_sat_i[k],
            ab, abi, epsr, epsi_tot, w.T_atm[k], c.t_prev[k], dqsdt, dt, t.qr2qv_evap_tend, t.nr_evap_tend);
        if (do_ice_production) {
          P::ice_deposition_sublimation(w.qi_incld[k], w.ni_incld[k], w.T_atm[k], w.qv_sat_l[k], w.qv_sat_i[k],
              epsi, abi, c.qv[k], inv_dt, t.qv2qi_vapdep_tend, t.qi2qv_sublim_tend, t.ni_sublim_tend, t.qc2qi_berg_tend);
        }
      }

      if (do_ice_production) {
        P::ice_nucleation(w.T_atm[k], w.inv_rho[k], c.ni[k], c.ni_activated[k], w.qv_supersat_i[k], inv_dt,
            predictNc, do_prescribed_CCN, t.qv2qi_nucleat_tend, t.ni_nucleat_tend, ro);
      }
      P::cloud_water_autoconversion(w.rho[k], w.qc_incld[k], w.nc_incld[k], c.inv_qc_relvar[k],
          t.qc2qr_autoconv_tend, t.nc2nr_autoconv_tend, t.ncautr, ro);
      P::droplet_self_collection(w.rho[k], w.inv_rho[k], w.qc_incld[k], w.mu_c[k], w.nu[k],
          t.nc2nr_autoconv_tend, t.nc_selfcollect_tend);
      P::cloud_rain_accretion(w.rho[k], w.inv_rho[k], w.qc_incld[k], w.nc_incld[k], w.qr_incld[k],
          c.inv_qc_relvar[k], t.qc2qr_accret_tend, t.nc_accret_tend, ro);
      P::rain_self_collection(w.rho[k], w.qr_incld[k], w.nr_incld[k], t.nr_selfcollect_tend, ro);

      P::back_to_cell_average(c.cld_frac_l[k], c.cld_frac_r[k], c.cld_frac_i[k], t);

      P::cloud_water_conservation(c.qc[k], dt, t);
      P::rain_water_conservation(c.qr[k], dt, t);
      P::ice_water_conservation(c.qi[k], dt, t);
      P::nc_conservation(c.nc[k], dt, t);
      P::nr_conservation(c.nr[k], dt, nmltratio, t);
      P::ni_conservation(c.ni[k], dt, t);
      P::ice_supersat_conservation(t.qv2qi_vapdep_tend, t.qv2qi_nucleat_tend, c.cld_frac_i[k], c.qv[k],
          w.qv_sat_i[k], c.th[k]/c.inv_exner[k], dt, t.qi2qv_sublim_tend, t.qr2qv_evap_tend);
      P::prevent_liq_supersaturation(c.pres[k], w.T_atm[k], c.qv[k], dt, t.qv2qi_vapdep_tend,
          t.qv2qi_nucleat_tend, t.qi2qv_sublim_tend, t.qr2qv_evap_tend, err);

      P::update_prognostic_ice(t, c.inv_exner[k], predictNc, wetgrowth, dt, nmltratio, rho_qm_cloud,
          c.th[k], c.qv[k], c.qi[k], c.ni[k], c.qm[k], c.bm[k], c.qc[k], c.nc[k], c.qr[k], c.nr[k]);
      P::update_prognostic_liquid(t, predictNc, do_prescribed_CCN, w.inv_rho[k], c.inv_exner[k], dt,
          c.th[k], c.qv[k], c.qc[k], c.nc[k], c.qr[k], c.nr[k]);

      c.qv2qi_depos_tend[k]  = t.qv2qi_vapdep_tend - t.qi2qv_sublim_tend + t.qv2qi_nucleat_tend;
      c.precip_total_tend[k] = t.qc2qr_accret_tend + t.qc2qr_autoconv_tend + t.qc2qr_ice_shed_tend + t.qc2qi_collect_tend;
      c.nevapr[k]            = t.qi2qv_sublim_tend + t.qr2qv_evap_tend;
      w.qr_evap_tend[k]      = t.qr2qv_evap_tend;
      c.vap_ice_exchange[k]  = t.qv2qi_vapdep_tend - t.qi2qv_sublim_tend + t.qv2qi_nucleat_tend;
      c.vap_liq_exchange[k]  = -t.qr2qv_evap_tend;
      c.liq_ice_exchange[k]  = t.qc2qi_hetero_freeze_tend + t.qr2qi_immers_freeze_tend - t.qi2qr_melt_tend +
                               t.qc2qi_berg_tend + t.qc2qi_collect_tend + t.qr2qi_collect_tend;

      if (c.qc[k] < qsmall) {
        c.qv[k] = c.qv[k] + c.qc[k];
        c.th[k] = c.th[k] - c.inv_exner[k] * c.qc[k] * latvap * inv_cp;
        c.qc[k] = 0; c.nc[k] = 0;
      } else hydrometeorsPresent = true;
      if (c.qr[k] < qsmall) {
        c.qv[k] = c.qv[k] + c.qr[k];
        c.th[k] = c.th[k] - c.inv_exner[k] * c.qr[k] * latvap * inv_cp;
        c.qr[k] = 0; c.nr[k] = 0;
      } else hydrometeorsPresent = true;
      if (c.qi[k] < qsmall) {
        c.qv[k] = c.qv[k] + c.qi[k];
        c.th[k] = c.th[k] - c.inv_exner[k] * c.qi[k] * (latvap+latice) * inv_cp;
        c.qi[k] = 0; c.ni[k] = 0; c.qm[k] = 0; c.bm[k] = 0;
      } else hydrometeorsPresent = true;

      w.pratot[k] = t.qc2qr_accret_tend;
      w.prctot[k] = t.qc2qr_autoconv_tend;

      w.ni_incld[k] = c.ni[k]/c.cld_frac_i[k];
      P::impose_max_total_ni(w.ni_incld[k], max_total_ni, w.inv_rho[k]);
      c.ni[k] = w.ni_incld[k]*c.cld_frac_i[k];

      P::calculate_incloud_mixingratios(
        c.qc[k], c.qr[k], c.qi[k], c.qm[k], c.nc[k], c.nr[k], c.ni[k], c.bm[k],
        w.inv_cld_frac_l[k], w.inv_cld_frac_i[k], w.inv_cld_frac_r[k],
        w.qc_incld[k], w.qr_incld[k], w.qi_incld[k], w.qm_incld[k], w.nc_incld[k], w.nr_incld[k],
        w.ni_incld[k], w.bm_incld[k]);
    }
  }

  // p3_cloud_sed_impl.hpp:17-147
  static void cloud_sedimentation(int nk, int ktop, int kbot, int kdir, Real dt, Real inv_dt, bool do_predict_nc,
      const Col& c, WS& w, int* diag, int* err) {
    Real* fluxes[2] = {w.flux[0], w.flux[1]};
    Real* vs[2]     = {w.V_q, w.V_n};
    Real* qnr[2]    = {c.qc, c.nc};
    const Real qsmall = C::QSMALL, bcn = C::bcn;
    bool log_qxpresent;
    const int k_qxtop = find_top(c.qc, qsmall, kbot, ktop, kdir, log_qxpresent);
    int nsub = 0, k_qxbot = -1;
    if (log_qxpresent) {
      Real dt_left = dt, prt_accum = 0.0;
      k_qxbot = find_bottom(c.qc, qsmall, kbot, k_qxtop, kdir, log_qxpresent);
      while (dt_left > C::dt_left_tol) {
        Real Co_max = 0.0;
        const int kmin = (kdir == 1 ? k_qxbot : k_qxtop), kmax = (kdir == 1 ? k_qxtop : k_qxbot);
        for (int k = 0; k < nk; ++k) { w.V_q[k] = 0; if (do_predict_nc) w.V_n[k] = 0; }
        for (int k = kmin; k <= kmax; ++k) {
          const bool qc_gt_small = w.qc_incld[k] > qsmall;
          if (qc_gt_small) {
            Real nu, cdist, cdist1;
            P::get_cloud_dsd2(w.qc_incld[k], w.nc_incld[k], w.mu_c[k], w.rho[k], nu, w.lamc[k], cdist, cdist1);
            c.nc[k] = w.nc_incld[k]*c.cld_frac_l[k];
            const Real dum = 1 / pm::pow(w.lamc[k], bcn);
            w.V_q[k] = w.acn[k]*pm::tgamma(4 + bcn + w.mu_c[k]) * dum / pm::tgamma(w.mu_c[k]+4);
            if (do_predict_nc)
              w.V_n[k] = w.acn[k]*pm::tgamma(1 + bcn + w.mu_c[k]) * dum / pm::tgamma(w.mu_c[k]+1);
            const Real Co_max_local = w.V_q[k] * dt_left * w.inv_dz[k];
            if (Co_max_local > Co_max) Co_max = Co_max_local;
          }
        }
        if (do_predict_nc)
          generalized_sedimentation<2>(w.rho, w.inv_rho, w.inv_dz, nk, k_qxtop, k_qxbot, kbot, kdir, Co_max, dt_left, prt_accum, fluxes, vs, qnr, err);
        else
          generalized_sedimentation<1>(w.rho, w.inv_rho, w.inv_dz, nk, k_qxtop, k_qxbot, kbot, kdir, Co_max, dt_left, prt_accum, fluxes, vs, qnr, err);
        for (int k = 0; k < nk; ++k) {
          w.qc_incld[k] = c.qc[k]/c.cld_frac_l[k];
          w.nc_incld[k] = c.nc[k]/c.cld_frac_l[k];
        }
        ++nsub;
      }
      *c.precip_liq_surf = prt_accum * C::INV_RHO_H2O * inv_dt;
    }
    for (int k = 0; k < nk; ++k) {
      w.qtend_ignore[k] = (c.qc[k] - w.qtend_ignore[k]) * inv_dt;
      w.ntend_ignore[k] = (c.nc[k] - w.ntend_ignore[k]) * inv_dt;
    }
    if (diag) { diag[D_QC_KTOP] = log_qxpresent || nsub ? k_qxtop : -1; diag[D_QC_KBOT] = k_qxbot; diag[D_QC_NSUB] = nsub; }
  }

  // p3_rain_sed_impl.hpp:17-34
  static void compute_rain_fall_velocity(const Tables& tb, Real qr_incld, Real rhofacr,
      Real& nr_incld, Real& mu_r, Real& lamr, Real& V_qr, Real& V_nr, const Runtime& ro) {
    Table3<Real> table;
    P::get_rain_dsd2(qr_incld, nr_incld, mu_r, lamr, ro);
    P::lookup(mu_r, lamr, table);
    V_qr = P::apply_table(tb.vm_table_vals, table) * rhofacr;
    V_nr = P::apply_table(tb.vn_table_vals, table) * rhofacr;
  }

  // p3_rain_sed_impl.hpp:39-170
  static void rain_sedimentation(int nk, int ktop, int kbot, int kdir, Real dt, Real inv_dt,
      const Tables& tb, const Col& c, WS& w, const Runtime& ro, int* diag, int* err) {
    Real* fluxes[2] = {w.flux[0], w.flux[1]};
    Real* vs[2]     = {w.V_q, w.V_n};
    Real* qnr[2]    = {c.qr, c.nr};
    const Real qsmall = C::QSMALL;
    bool log_qxpresent;
    const int k_qxtop = find_top(c.qr, qsmall, kbot, ktop, kdir, log_qxpresent);
    int nsub = 0, k_qxbot = -1;
    if (log_qxpresent) {
      Real dt_left = dt, prt_accum = 0.0;
      k_qxbot = find_bottom(c.qr, qsmall, kbot, k_qxtop, kdir, log_qxpresent);
      while (dt_left > C::dt_left_tol) {
        Real Co_max = 0.0;
        int kmin = (kdir == 1 ? k_qxbot : k_qxtop), kmax = (kdir == 1 ? k_qxtop : k_qxbot);
        for (int k = 0; k < nk; ++k) { w.V_q[k] = 0; w.V_n[k] = 0; }
        for (int k = kmin; k <= kmax; ++k) {
          const bool qr_gt_small = w.qr_incld[k] > qsmall;
          if (qr_gt_small) {
            compute_rain_fall_velocity(tb, w.qr_incld[k], w.rhofacr[k], w.nr_incld[k], w.mu_r[k], w.lamr[k],
                                       w.V_q[k], w.V_n[k], ro);
            c.nr[k] = w.nr_incld[k]*c.cld_frac_r[k];
            const Real Co_max_local = w.V_q[k] * dt_left * w.inv_dz[k];
            if (Co_max_local > Co_max) Co_max = Co_max_local;
          }
        }
        generalized_sedimentation<2>(w.rho, w.inv_rho, w.inv_dz, nk, k_qxtop, k_qxbot, kbot, kdir, Co_max, dt_left, prt_accum, fluxes, vs, qnr, err);
        for (int k = 0; k < nk; ++k) {
          w.qr_incld[k] = c.qr[k]/c.cld_frac_r[k];
          w.nr_incld[k] = c.nr[k]/c.cld_frac_r[k];
        }
        // precip_liq_flux output (l.139-153): level interface k+1 accumulates flux_qx(k)
        kmin = (kdir == 1 ? k_qxbot+1 : k_qxtop+1);
        kmax = (kdir == 1 ? k_qxtop+1 : k_qxbot+1);
        for (int k = kmin; k <= kmax; ++k) {
          int idx = k-1; if (idx < 0) idx = 0;
          c.precip_liq_flux[k] = c.precip_liq_flux[k] + w.flux[0][idx];
        }
        ++nsub;
      }
      *c.precip_liq_surf += prt_accum * C::INV_RHO_H2O * inv_dt;
    }
    for (int k = 0; k < nk; ++k) {
      w.qtend_ignore[k] = (c.qr[k] - w.qtend_ignore[k]) * inv_dt;
      w.ntend_ignore[k] = (c.nr[k] - w.ntend_ignore[k]) * inv_dt;
    }
    if (diag) { diag[D_QR_KTOP] = log_qxpresent || nsub ? k_qxtop : -1; diag[D_QR_KBOT] = k_qxbot; diag[D_QR_NSUB] = nsub; }
  }

  // p3_ice_sed_impl.hpp:69-205
  static void ice_sedimentation(int nk, int ktop, int kbot, int kdir, Real dt, Real inv_dt,
      const Tables& tb, const Col& c, WS& w, const Runtime& ro, int* diag, int* err) {
    Real* fluxes[4] = {w.flux[0], w.flux[1], w.flux[2], w.flux[3]};   // qit, nit, qir, bir
    Real* vs[4]     = {w.V_q, w.V_n, w.V_q, w.V_q};
    Real* qnr[4]    = {c.qi, c.ni, c.qm, c.bm};
    const Real qsmall = C::QSMALL, nsmall = C::NSMALL;
    const Real ice_sedimentation_factor = ro.ice_sedimentation_factor;
    bool log_qxpresent;
    const int k_qxtop = find_top(c.qi, qsmall, kbot, ktop, kdir, log_qxpresent);
    int nsub = 0, k_qxbot = -1;
    if (log_qxpresent) {
      Real dt_left = dt, prt_accum = 0.0;
      k_qxbot = find_bottom(c.qi, qsmall, kbot, k_qxtop, kdir, log_qxpresent);
      while (dt_left > C::dt_left_tol) {
        Real Co_max = 0.0;
        const int kmin = (kdir == 1 ? k_qxbot : k_qxtop), kmax = (kdir == 1 ? k_qxtop : k_qxbot);
        for (int k = 0; k < nk; ++k) { w.V_q[k] = 0; w.V_n[k] = 0; }
        for (int k = kmin; k <= kmax; ++k) {
          const bool qi_gt_small = w.qi_incld[k] > qsmall;
          if (qi_gt_small) {
            w.ni_incld[k] = P::max(w.ni_incld[k], nsmall);
            const Real rhop = P::calc_bulk_rho_rime(w.qi_incld[k], w.qm_incld[k], w.bm_incld[k], ro);
            c.qm[k] = w.qm_incld[k]*c.cld_frac_i[k];
            c.bm[k] = w.bm_incld[k]*c.cld_frac_i[k];
            TableIce<Real> tab;
            P::lookup_ice(w.qi_incld[k], w.ni_incld[k], w.qm_incld[k], rhop, tab);
            const Real table_val_ni_fallspd = P::apply_table_ice(0, tb, tab);
            const Real table_val_qi_fallspd = P::apply_table_ice(1, tb, tab);
            const Real table_val_ni_lammax  = P::apply_table_ice(6, tb, tab);
            const Real table_val_ni_lammin  = P::apply_table_ice(7, tb, tab);
            w.ni_incld[k] = P::min(w.ni_incld[k], table_val_ni_lammax * w.ni_incld[k]);
            w.ni_incld[k] = P::max(w.ni_incld[k], table_val_ni_lammin * w.ni_incld[k]);
            c.ni[k] = w.ni_incld[k] * c.cld_frac_i[k];
            w.V_q[k] = ice_sedimentation_factor * table_val_qi_fallspd * w.rhofaci[k];
            w.V_n[k] = ice_sedimentation_factor * table_val_ni_fallspd * w.rhofaci[k];
            const Real Co_max_local = w.V_q[k] * dt_left * w.inv_dz[k];
            if (Co_max_local > Co_max) Co_max = Co_max_local;
          }
        }
        generalized_sedimentation<4>(w.rho, w.inv_rho, w.inv_dz, nk, k_qxtop, k_qxbot, kbot, kdir, Co_max, dt_left, prt_accum, fluxes, vs, qnr, err);
        for (int k = 0; k < nk; ++k) {
          w.qi_incld[k] = c.qi[k]/c.cld_frac_i[k];
          w.ni_incld[k] = c.ni[k]/c.cld_frac_i[k];
          w.qm_incld[k] = c.qm[k]/c.cld_frac_i[k];
          w.bm_incld[k] = c.bm[k]/c.cld_frac_i[k];
        }
        ++nsub;
      }
      *c.precip_ice_surf += prt_accum * C::INV_RHO_H2O * inv_dt;
    }
    for (int k = 0; k < nk; ++k) {
      w.qtend_ignore[k] = (c.qi[k] - w.qtend_ignore[k]) * inv_dt;
      w.ntend_ignore[k] = (c.ni[k] - w.ntend_ignore[k]) * inv_dt;
    }
    if (diag) { diag[D_QI_KTOP] = log_qxpresent || nsub ? k_qxtop : -1; diag[D_QI_KBOT] = k_qxbot; diag[D_QI_NSUB] = nsub; }
  }

  // p3_ice_sed_impl.hpp:210-270
  static void homogeneous_freezing(int /*nk*/, int ktop, int kbot, int kdir, const Col& c, WS& w) {
    const Real qsmall = C::QSMALL, nsmall = C::NSMALL, T_homogfrz = C::T_homogfrz;
    const Real inv_rho_rimeMax = C::INV_RHO_RIMEMAX, inv_cp = C::INV_CP, latice = C::LatIce;
    const int kmin = (kdir == 1 ? kbot : ktop), kmax = (kdir == 1 ? ktop : kbot);
    for (int k = kmin; k <= kmax; ++k) {
      const bool t_lt_homogf = w.T_atm[k] < T_homogfrz;
      const bool qc_ge_small = t_lt_homogf && c.qc[k] >= qsmall;
      const bool qr_ge_small = t_lt_homogf && c.qr[k] >= qsmall;
      const Real Qc_nuc = c.qc[k], Qr_nuc = c.qr[k];
      const Real Nc_nuc = P::max(c.nc[k], nsmall), Nr_nuc = P::max(c.nr[k], nsmall);
      if (qc_ge_small) {
        c.qm[k] = c.qm[k] + Qc_nuc;
        c.qi[k] = c.qi[k] + Qc_nuc;
        c.bm[k] = c.bm[k] + Qc_nuc*inv_rho_rimeMax;
        c.ni[k] = c.ni[k] + Nc_nuc;
        c.th[k] = c.th[k] + c.inv_exner[k]*Qc_nuc*latice*inv_cp;
      }
      if (qr_ge_small) {
        c.qm[k] = c.qm[k] + Qr_nuc;
        c.qi[k] = c.qi[k] + Qr_nuc;
        c.bm[k] = c.bm[k] + Qr_nuc*inv_rho_rimeMax;
        c.ni[k] = c.ni[k] + Nr_nuc;
        c.th[k] = c.th[k] + c.inv_exner[k]*Qr_nuc*latice*inv_cp;
      }
      if (qc_ge_small) { c.qc[k] = 0; c.nc[k] = 0; }
      if (qr_ge_small) { c.qr[k] = 0; c.nr[k] = 0; }
    }
  }

  // p3_main_impl_part3.hpp:21-204
  static void p3_main_part3(int nk, const Tables& tb, const Col& c, WS& w, const Runtime& ro) {
    const Real qsmall = C::QSMALL, inv_cp = C::INV_CP, nsmall = C::NSMALL, latvap = C::LatVap, latice = C::LatIce;
    const Real max_total_ni = ro.max_total_ni;
    for (int k = 0; k < nk; ++k) {
      Real ignore1 = 0, ignore2 = 0;
      Real table_val_qi_fallspd = 0, table_val_ice_eff_radius = 0, table_val_ni_lammax = 0, table_val_ni_lammin = 0,
           table_val_ice_reflectivity = 0, table_val_ice_mean_diam = 0, table_val_ice_bulk_dens = 0;
      // Cloud
      {
        const bool qc_gt_small = c.qc[k] >= qsmall;
        const Real qc_incld = c.qc[k]/c.cld_frac_l[k];
        Real nc_incld = c.nc[k]/c.cld_frac_l[k];
        if (qc_gt_small) {
          P::get_cloud_dsd2(qc_incld, nc_incld, w.mu_c[k], w.rho[k], w.nu[k], w.lamc[k], ignore1, ignore2);
          c.nc[k] = nc_incld*c.cld_frac_l[k];
          c.diag_eff_radius_qc[k] = sp(0.5) * (w.mu_c[k] + 3) / w.lamc[k];
        } else {
          c.qv[k] = c.qv[k]+c.qc[k];
          c.th[k] = c.th[k]-c.inv_exner[k]*c.qc[k]*latvap*inv_cp;
          c.vap_liq_exchange[k] = c.vap_liq_exchange[k] - c.qc[k];
          c.qc[k] = 0; c.nc[k] = 0;
        }
      }
      // Rain
      {
        const bool qr_gt_small = c.qr[k] >= qsmall;
        const Real qr_incld = c.qr[k]/c.cld_frac_r[k];
        Real nr_incld = c.nr[k]/c.cld_frac_r[k];
        if (qr_gt_small) {
          P::get_rain_dsd2(qr_incld, nr_incld, w.mu_r[k], w.lamr[k], ro);
          c.nr[k] = nr_incld*c.cld_frac_r[k];
          w.ze_rain[k] = c.nr[k]*(w.mu_r[k]+6)*(w.mu_r[k]+5)*(w.mu_r[k]+4)*
                         (w.mu_r[k]+3)*(w.mu_r[k]+2)*(w.mu_r[k]+1)/pm::pow(w.lamr[k], sp(6.0));
          w.ze_rain[k] = P::max(w.ze_rain[k], sp(1.e-22));
          c.diag_eff_radius_qr[k] = sp(1.5) / w.lamr[k];
        } else {
          c.qv[k] = c.qv[k] + c.qr[k];
          c.th[k] = c.th[k] - c.inv_exner[k]*c.qr[k]*latvap*inv_cp;
          c.vap_liq_exchange[k] = c.vap_liq_exchange[k] - c.qr[k];
          c.qr[k] = 0; c.nr[k] = 0;
        }
      }
      // Ice
      {
        const bool qi_gt_small = c.qi[k] >= qsmall;
        c.ni[k] = P::max(c.ni[k], nsmall);
        Real qi_incld = c.qi[k]/c.cld_frac_i[k];
        Real ni_incld = c.ni[k]/c.cld_frac_i[k];
        Real qm_incld = c.qm[k]/c.cld_frac_i[k];
        Real bm_incld = c.bm[k]/c.cld_frac_i[k];
        Real rhop = 0;
        if (qi_gt_small) {
          rhop = P::calc_bulk_rho_rime(qi_incld, qm_incld, bm_incld, ro);
          c.qm[k] = qm_incld*c.cld_frac_i[k];
          c.bm[k] = bm_incld*c.cld_frac_i[k];
        }
        P::impose_max_total_ni(ni_incld, max_total_ni, w.inv_rho[k]);
        if (qi_gt_small) {
          TableIce<Real> table_ice;
          P::lookup_ice(qi_incld, ni_incld, qm_incld, rhop, table_ice);
          table_val_qi_fallspd       = P::apply_table_ice(1,  tb, table_ice);
          table_val_ice_eff_radius   = P::apply_table_ice(5,  tb, table_ice);
          table_val_ni_lammax        = P::apply_table_ice(6,  tb, table_ice);
          table_val_ni_lammin        = P::apply_table_ice(7,  tb, table_ice);
          table_val_ice_reflectivity = P::apply_table_ice(8,  tb, table_ice);
          table_val_ice_mean_diam    = P::apply_table_ice(10, tb, table_ice);
          table_val_ice_bulk_dens    = P::apply_table_ice(11, tb, table_ice);
          ni_incld = P::min(ni_incld, table_val_ni_lammax * ni_incld);
          ni_incld = P::max(ni_incld, table_val_ni_lammin * ni_incld);
        }
        c.ni[k] = ni_incld*c.cld_frac_i[k];
        if (qi_gt_small) {
          if (c.qm[k] < qsmall) { c.qm[k] = 0; c.bm[k] = 0; }
          w.diag_vm_qi[k] = table_val_qi_fallspd * w.rhofaci[k];
          c.diag_eff_radius_qi[k] = table_val_ice_eff_radius;
          w.diag_diam_qi[k] = table_val_ice_mean_diam;
          c.rho_qi[k] = table_val_ice_bulk_dens;
          w.ze_ice[k] = w.ze_ice[k] + sp(0.1892)*table_val_ice_reflectivity*ni_incld*w.rho[k];
          w.ze_ice[k] = P::max(w.ze_ice[k], sp(1.e-22));
          w.ze_ice[k] = w.ze_ice[k]*c.cld_frac_i[k];
        } else {
          c.qv[k] = c.qv[k] + c.qi[k];
          c.th[k] = c.th[k] - c.inv_exner[k]*c.qi[k]*(latvap+latice)*inv_cp;
          c.qi[k] = 0; c.ni[k] = 0; c.qm[k] = 0; c.bm[k] = 0;
          w.diag_diam_qi[k] = 0;
        }
      }
      w.diag_equiv_reflectivity[k] = 10 * pm::log10((w.ze_rain[k] + w.ze_ice[k])*sp(1.e18));
      if (c.qr[k] < qsmall) c.nr[k] = 0;
    }
  }

  // p3_main_impl.hpp:76-331, body of the "p3 main loop" for one column i.
  // diag: optional D_NUM ints for this column.
  static void p3_main_column(const Runtime& ro, const Col& c, Real dt, bool predictNc, bool prescribedCCN,
                             const Tables& tb, int nk, WS& w, int* diag) {
    const Real inv_dt = 1 / dt;
    const int kdir = -1, ktop = 0, kbot = nk-1;
    int err = 0;
    int dloc[D_NUM];
    for (int j = 0; j < D_NUM; ++j) dloc[j] = -1;
    dloc[D_QC_NSUB] = dloc[D_QR_NSUB] = dloc[D_QI_NSUB] = 0;
    bool nucleationPossible, hydrometeorsPresent;

    p3_main_init(nk, c, w);
    p3_main_part1(nk, predictNc, prescribedCCN, dt, c, w, nucleationPossible, hydrometeorsPresent, ro, &err);
    dloc[D_NUC_POSSIBLE] = nucleationPossible; dloc[D_HYDRO_PART1] = hydrometeorsPresent;
    if (nucleationPossible || hydrometeorsPresent) {
      p3_main_part2(nk, predictNc, prescribedCCN, dt, inv_dt, tb, c, w, hydrometeorsPresent, ro, &err);
      dloc[D_HYDRO_PART2] = hydrometeorsPresent;
      if (hydrometeorsPresent) {
        cloud_sedimentation(nk, ktop, kbot, kdir, dt, inv_dt, predictNc, c, w, dloc, &err);
        rain_sedimentation(nk, ktop, kbot, kdir, dt, inv_dt, tb, c, w, ro, dloc, &err);
        ice_sedimentation(nk, ktop, kbot, kdir, dt, inv_dt, tb, c, w, ro, dloc, &err);
        if (ro.do_ice_production) homogeneous_freezing(nk, ktop, kbot, kdir, c, w);
        p3_main_part3(nk, tb, c, w, ro);
      }
    }
    dloc[D_ERROR] = err;
    if (diag) for (int j = 0; j < D_NUM; ++j) diag[j] = dloc[j];
  }
};

} // namespace p3o
#endif
