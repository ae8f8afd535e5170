// ORACLE — TEST INFRASTRUCTURE ONLY (see p3_oracle_types.hpp header).
// Per-level process functions of P3, restated with pack-size-1 (scalar) semantics:
//   x.set(m, v)  ->  if (m) x = v;      if (m.any()) {..} -> if (m) {..}
//   IntSmallPack(realpack) -> (int) truncation toward zero (ekat_pack.hpp:192-194)
// Citations: p3/impl/<file>:<lines> in /root/reference/components/eamxx/src/physics.
#ifndef P3_ORACLE_PROCS_HPP
#define P3_ORACLE_PROCS_HPP

#include "p3_oracle_types.hpp"

namespace p3o {

template <typename Real>
struct Procs {
  using C = Constants<Real>;
  using Runtime = P3Runtime<Real>;
  using Tables  = P3LookupTables<Real>;
  static Real sp(double x) { return static_cast<Real>(x); }
  static Real square(Real x) { return x*x; }
  static Real cube(Real x) { return x*x*x; }
  static Real max(Real a, Real b) { return a < b ? b : a; }   // ekat::impl::max semantics
  static Real min(Real a, Real b) { return a < b ? a : b; }
  static int  imax(int a, int b) { return a < b ? b : a; }
  static int  imin(int a, int b) { return a < b ? a : b; }

  // share/physics_saturation_impl.hpp:17-37 — returns error bits instead of aborting.
  static int check_temperature(Real t_atm) {
    int err = 0;
    if (t_atm <= 0) err |= 1;
    if (std::isnan(t_atm)) err |= 2;
    return err;
  }

  // share/physics_saturation_impl.hpp:41-83
  static Real MurphyKoop_svp(Real t_atm, bool ice, int* err) {
    if (err) *err |= check_temperature(t_atm);
    Real result = 0;
    const Real tmelt = C::Tmelt;
    const bool ice_mask = (t_atm < tmelt) && ice;
    if (ice_mask) {
      const Real ic[] = {sp(9.550426), sp(5723.265), sp(3.53068), sp(0.00728332)};
      result = pm::exp(ic[0] - (ic[1] / t_atm) + (ic[2] * pm::log(t_atm)) - (ic[3] * t_atm));
    } else {
      const Real lq[] = {sp(54.842763), sp(6763.22), sp(4.210), sp(0.000367), sp(0.0415), sp(218.8),
                         sp(53.878), sp(1331.22), sp(9.44523), sp(0.014025)};
      const Real logt = pm::log(t_atm);
      result = pm::exp(lq[0] - (lq[1] / t_atm) - (lq[2] * logt) + (lq[3] * t_atm) +
                        (pm::tanh(lq[4] * (t_atm - lq[5])) * (lq[6] - (lq[7] / t_atm) -
                                                             (lq[8] * logt) + lq[9] * t_atm)));
    }
    return result;
  }

  // share/physics_saturation_impl.hpp:130-155 (func_idx = MurphyKoop, the only one P3 uses)
  static Real qv_sat_dry(Real t_atm, Real p_atm_dry, bool ice, int* err) {
    const Real e_pres = MurphyKoop_svp(t_atm, ice, err);
    return C::ep_2 * e_pres / max(p_atm_dry, sp(1.e-3));
  }

  // p3_incloud_mixingratios_impl.hpp:12-69 (note bm_incld uses inv_cld_frac_l, l.49)
  static void calculate_incloud_mixingratios(
      Real qc, Real qr, Real qi, Real qm, Real nc, Real nr, Real ni, Real bm,
      Real inv_cld_frac_l, Real inv_cld_frac_i, Real inv_cld_frac_r,
      Real& qc_incld, Real& qr_incld, Real& qi_incld, Real& qm_incld,
      Real& nc_incld, Real& nr_incld, Real& ni_incld, Real& bm_incld) {
    const Real qsmall = C::QSMALL, incloud_limit = C::incloud_limit, precip_limit = C::precip_limit;
    if (qc >= qsmall) { qc_incld = qc*inv_cld_frac_l; nc_incld = max(nc*inv_cld_frac_l, 0); }
    else              { qc_incld = 0; nc_incld = 0; }
    if (qi >= qsmall) { qi_incld = qi*inv_cld_frac_i; ni_incld = max(ni*inv_cld_frac_i, 0); }
    else              { qi_incld = 0; ni_incld = 0; }
    if (qi >= qsmall && qm >= qsmall) { qm_incld = qm*inv_cld_frac_i; bm_incld = max(bm*inv_cld_frac_l, 0); }
    else                              { qm_incld = 0; bm_incld = 0; }
    if (qr >= qsmall) { qr_incld = qr*inv_cld_frac_r; nr_incld = max(nr*inv_cld_frac_r, 0); }
    else              { qr_incld = 0; nr_incld = 0; }
    if (qc_incld > incloud_limit || qi_incld > incloud_limit || qr_incld > precip_limit ||
        bm_incld > incloud_limit) {
      qc_incld = min(qc_incld, incloud_limit);
      qi_incld = min(qi_incld, incloud_limit);
      bm_incld = min(bm_incld, incloud_limit);
      qr_incld = min(qr_incld, precip_limit);
    }
  }

  // p3_get_time_space_phys_variables_impl.hpp:13-51
  static void get_time_space_phys_variables(
      Real T_atm, Real pres, Real rho, Real qv_sat_l, Real qv_sat_i,
      Real& mu, Real& dv, Real& sc, Real& dqsdt, Real& dqsidt, Real& ab, Real& abi, Real& kap, Real& eii) {
    mu = sp(1.496e-6) * pm::pow(T_atm, sp(1.5))/(T_atm+120);
    dv = sp(8.794e-5) * pm::pow(T_atm, sp(1.81))/pres;
    sc = mu/(rho*dv);
    const Real RV = C::RV, INV_CP = C::INV_CP, latvap = C::LatVap, latice = C::LatIce;
    const Real tval1 = 253.15, tval2 = 273.15, dtval = 20;
    const Real dum = 1/(RV*square(T_atm));
    dqsdt  = latvap*qv_sat_l*dum;
    dqsidt = (latvap+latice)*qv_sat_i*dum;
    ab  = 1+dqsdt*latvap*INV_CP;
    abi = 1+dqsidt*(latvap+latice)*INV_CP;
    kap = sp(1.414e+3)*mu;
    if (T_atm < tval1)      eii = sp(0.001);
    else if (T_atm < tval2) eii = sp(0.001)+( T_atm-sp(tval1) )*( sp(0.3) - sp(0.001) )/dtval;
    else                    eii = sp(0.3);
  }

  // p3_dsd2_impl.hpp:16-73 (iparam == 3, so the dnu interpolation branch l.49-54 is dead)
  static void get_cloud_dsd2(Real qc, Real& nc, Real& mu_c, Real rho, Real& nu,
                             Real& lamc, Real& cdist, Real& cdist1) {
    lamc = 0; cdist = 0; cdist1 = 0; nu = 0; mu_c = 0;
    const Real qsmall = C::QSMALL;
    if (qc >= qsmall) {
      const Real nsmall = C::NSMALL, cons1 = C::CONS1;
      nc = max(nc, nsmall);
      Real mu_c_local = sp(0.0005714)*(nc * sp(1.e-6) * rho) + sp(0.2714);
      mu_c_local = 1/(mu_c_local * mu_c_local) - 1;
      mu_c_local = max(mu_c_local, 2);
      mu_c_local = min(mu_c_local, 15);
      mu_c = mu_c_local;
      lamc = pm::cbrt(cons1 * nc * (mu_c + 3) * (mu_c + 2) * (mu_c + 1) / qc);
      const Real lammin = (mu_c + 1)*sp(2.5e+4);
      const Real lammax = (mu_c + 1)*sp(1.e+6);
      const bool lamc_lt_min = lamc < lammin, lamc_gt_max = lamc > lammax;
      if (lamc_lt_min) lamc = lammin;
      if (lamc_gt_max) lamc = lammax;
      if (lamc_lt_min || lamc_gt_max)
        nc = 6 * (lamc * lamc * lamc) * qc / (C::Pi * C::RHO_H2O * (mu_c + 3) * (mu_c + 2) * (mu_c + 1));
      cdist  = nc * (mu_c+1) / lamc;
      cdist1 = nc / pm::tgamma(mu_c + 1);
    }
  }

  // p3_dsd2_impl.hpp:76-127
  static void get_rain_dsd2(Real qr, Real& nr, Real& mu_r, Real& lamr, const Runtime& ro) {
    const Real nsmall = C::NSMALL, qsmall = C::QSMALL, cons1 = C::CONS1;
    lamr = 0;
    const Real mu_r_const = ro.constant_mu_rain;
    if (qr >= qsmall) {
      const Real nr_lim = max(nr, nsmall);
      mu_r = mu_r_const;
      const Real mass_to_d3_factor = cons1 * (mu_r + 3) * (mu_r + 2) * (mu_r + 1);
      lamr = pm::cbrt(mass_to_d3_factor * nr_lim / qr);
      const Real lammax = (mu_r+sp(1.))*sp(1.e+5);
      const Real lammin = (mu_r+sp(1.))*500;
      const bool lt = lamr < lammin, gt = lamr > lammax;
      nr = nr_lim;
      if (lt || gt) {
        if (lt) lamr = lammin;
        if (gt) lamr = lammax;
        nr = lamr*lamr*lamr * qr / mass_to_d3_factor;
      }
    }
  }

  // p3_dsd2_impl.hpp:130-149
  static void get_cdistr_logn0r(Real qr, Real nr, Real mu_r, Real lamr, Real& cdistr, Real& logn0r) {
    cdistr = 0; logn0r = 0;
    if (qr >= C::QSMALL) {
      cdistr = nr/pm::tgamma(mu_r + 1);
      logn0r = pm::log10(cdistr) + (mu_r + 1) * pm::log10(lamr);
    }
  }

  // p3_impose_max_total_ni_impl.hpp:12-27
  static void impose_max_total_ni(Real& ni_local, Real max_total_ni, Real inv_rho_local) {
    if (ni_local >= sp(1e-20))
      ni_local = ni_local*min(max_total_ni * inv_rho_local/ni_local, 1);
  }

  // p3_ice_sed_impl.hpp:18-64
  static Real calc_bulk_rho_rime(Real qi_tot, Real& qi_rim, Real& bi_rim, const Runtime& ro) {
    const Real bsmall = C::BSMALL, qsmall = C::QSMALL;
    const Real min_rime_rho = ro.min_rime_rho, max_rime_rho = ro.max_rime_rho;
    Real rho_rime = 0;
    const bool bi_rim_gt_small = bi_rim >= bsmall;
    const bool bi_rim_lt_small = bi_rim <  bsmall;
    if (bi_rim_gt_small) rho_rime = qi_rim / bi_rim;
    const bool rho_rime_lt_min = rho_rime < min_rime_rho;
    const bool rho_rime_gt_max = rho_rime > max_rime_rho;
    if (bi_rim_gt_small && rho_rime_lt_min) rho_rime = min_rime_rho;
    if (bi_rim_gt_small && rho_rime_gt_max) rho_rime = max_rime_rho;
    if (bi_rim_gt_small && (rho_rime_gt_max || rho_rime_lt_min)) bi_rim = qi_rim / rho_rime;
    if (bi_rim_lt_small) { qi_rim = 0; bi_rim = 0; rho_rime = 0; }
    if ((qi_rim > qi_tot) && (rho_rime > 0)) { qi_rim = qi_tot; bi_rim = qi_rim / rho_rime; }
    if (qi_rim < qsmall) { qi_rim = 0; bi_rim = 0; }
    return rho_rime;
  }

  // p3_table_ice_impl.hpp:83-138
  static void lookup_ice(Real qi, Real ni, Real qm, Real rhop, TableIce<Real>& tab) {
    const Real c = static_cast<Real>(P3C::lookup_table_1a_dum1_c);
    tab.dum1 = (pm::log10(qi/ni)+18) * c - 10;
    tab.dumi = static_cast<int>(tab.dum1);
    tab.dum1 = min(tab.dum1, static_cast<Real>(P3C::isize));
    tab.dum1 = max(tab.dum1, sp(1.));
    tab.dumi = imax(1, tab.dumi);
    tab.dumi = imin(P3C::isize-1, tab.dumi);
    tab.dum4  = (qm/qi)*3 + 1;
    tab.dumii = static_cast<int>(tab.dum4);
    tab.dum4  = min(tab.dum4, static_cast<Real>(P3C::rimsize));
    tab.dum4  = max(tab.dum4, sp(1.));
    tab.dumii = imax(1, tab.dumii);
    tab.dumii = imin(P3C::rimsize-1, tab.dumii);
    if (rhop <= 650) tab.dum5 = (rhop-50)*sp(0.005) + 1;
    else             tab.dum5 = (rhop-650)*sp(0.004) + 4;
    tab.dumjj = static_cast<int>(tab.dum5);
    tab.dum5  = min(tab.dum5, static_cast<Real>(P3C::densize));
    tab.dum5  = max(tab.dum5, sp(1.));
    tab.dumjj = imax(1, tab.dumjj);
    tab.dumjj = imin(P3C::densize-1, tab.dumjj);
    tab.dum6  = -99;
    tab.dumzz = -99;
    tab.dumi -= 1; tab.dumjj -= 1; tab.dumii -= 1; tab.dumzz -= 1;
  }

  // p3_table_ice_impl.hpp:143-173
  static void lookup_rain(Real qr, Real nr, TableRain<Real>& tab) {
    const Real qsmall = C::QSMALL;
    const bool gt_small = qr > qsmall && nr > 0.0;
    Real dumlr = 1;
    if (gt_small) dumlr = pm::cbrt(qr/(C::Pi * C::RHO_H2O * nr));
    tab.dum3 = (pm::log10(1*dumlr) + 5)*sp(10.70415);
    tab.dumj = static_cast<int>(tab.dum3);
    tab.dum3 = min(tab.dum3, static_cast<Real>(P3C::rcollsize));
    tab.dum3 = max(tab.dum3, 1);
    tab.dumj = imax(1, tab.dumj);
    tab.dumj = imin(P3C::rcollsize-1, tab.dumj);
    if (!gt_small) { tab.dumj = 1; tab.dum3 = 1; }
    tab.dumj -= 1;
  }

  // p3_table_ice_impl.hpp:178-217
  static Real apply_table_ice(int idx, const Tables& t, const TableIce<Real>& tab) {
    const auto& T = t.ice_table_vals;
    const int jj = tab.dumjj, ii = tab.dumii, i = tab.dumi;
    Real iproc1 = T[jj][ii][i][idx] + (tab.dum1-Real(i)-1) * (T[jj][ii][i+1][idx] - T[jj][ii][i][idx]);
    Real gproc1 = T[jj][ii+1][i][idx] + (tab.dum1-Real(i)-1) * (T[jj][ii+1][i+1][idx] - T[jj][ii+1][i][idx]);
    const Real tmp1 = iproc1 + (tab.dum4-Real(ii)-1) * (gproc1-iproc1);
    iproc1 = T[jj+1][ii][i][idx] + (tab.dum1-Real(i)-1) * (T[jj+1][ii][i+1][idx] - T[jj+1][ii][i][idx]);
    gproc1 = T[jj+1][ii+1][i][idx] + (tab.dum1-Real(i)-1) * (T[jj+1][ii+1][i+1][idx] - T[jj+1][ii+1][i][idx]);
    const Real tmp2 = iproc1+(tab.dum4 - Real(ii) - 1) * (gproc1-iproc1);
    return tmp1 + (tab.dum5 - Real(jj) - 1) * (tmp2-tmp1);
  }

  // p3_table_ice_impl.hpp:222-297
  static Real apply_table_coll(int idx, const Tables& t, const TableIce<Real>& ti, const TableRain<Real>& tr) {
    const auto& T = t.collect_table_vals;
    const int jj = ti.dumjj, ii = ti.dumii, i = ti.dumi, j = tr.dumj;
    const Real f1 = ti.dum1 - Real(i) - 1;
    const Real f3 = tr.dum3 - Real(j) - 1;
    Real dproc1 = T[jj][ii][i][j][idx] + f1 * (T[jj][ii][i+1][j][idx] - T[jj][ii][i][j][idx]);
    Real dproc2 = T[jj][ii][i][j+1][idx] + f1 * (T[jj][ii][i+1][j+1][idx] - T[jj][ii][i][j+1][idx]);
    Real iproc1 = dproc1 + f3 * (dproc2 - dproc1);
    dproc1 = T[jj][ii+1][i][j][idx] + f1 * (T[jj][ii+1][i+1][j][idx] - T[jj][ii+1][i][j][idx]);
    dproc2 = T[jj][ii+1][i][j+1][idx] + f1 * (T[jj][ii+1][i+1][j+1][idx] - T[jj][ii+1][i][j+1][idx]);
    Real gproc1 = dproc1 + f3 * (dproc2-dproc1);
    const Real tmp1 = iproc1 + (ti.dum4 - Real(ii) - 1) * (gproc1-iproc1);
    dproc1 = T[jj+1][ii][i][j][idx] + f1 * (T[jj+1][ii][i+1][j][idx] - T[jj+1][ii][i][j][idx]);
    dproc2 = T[jj+1][ii][i][j+1][idx] + f1 * (T[jj+1][ii][i+1][j+1][idx] - T[jj+1][ii][i][j+1][idx]);
    iproc1 = dproc1 + f3 * (dproc2-dproc1);
    dproc1 = T[jj+1][ii+1][i][j][idx] + f1 * (T[jj+1][ii+1][i+1][j][idx] - T[jj+1][ii+1][i][j][idx]);
    dproc2 = T[jj+1][ii+1][i][j+1][idx] + f1 * (T[jj+1][ii+1][i+1][j+1][idx] - T[jj+1][ii+1][i][j+1][idx]);
    gproc1 = dproc1 + f3 * (dproc2-dproc1);
    const Real tmp2 = iproc1 + (ti.dum4 - Real(ii) - 1) * (gproc1-iproc1);
    return tmp1 + (ti.dum5 - Real(jj) - 1) * (tmp2-tmp1);
  }

  // p3_table3_impl.hpp:17-65
  static void lookup(Real mu_r, Real lamr, Table3<Real>& tab) {
    const Real dum1 = (mu_r+1) / lamr;
    tab.dumii = 1;
    if (dum1 <= sp(195.e-6)) {
      const Real inv_dum3 = sp(0.1);
      Real rdumii = (dum1 * sp(1.e6) + 5) * inv_dum3;
      rdumii = max(rdumii, sp( 1.0));
      rdumii = min(rdumii, sp(20.0));
      int dumii = static_cast<int>(rdumii);
      dumii = imax(dumii,  1);
      dumii = imin(dumii, 20);
      tab.rdumii = rdumii; tab.dumii = dumii;
    } else {
      const Real inv_dum3 = C::THIRD * sp(0.1);
      Real rdumii = (dum1 * sp(1.e+6) - 195) * inv_dum3 + 20;
      rdumii = max(rdumii, sp( 20.0));
      rdumii = min(rdumii, sp(300.0));
      int dumii = static_cast<int>(rdumii);
      dumii = imax(dumii, 20);
      dumii = imin(dumii,299);
      tab.rdumii = rdumii; tab.dumii = dumii;
    }
    Real rdumjj = mu_r + 1;
    rdumjj = max(rdumjj, 1);
    rdumjj = min(rdumjj, 10);
    int dumjj = static_cast<int>(rdumjj);
    dumjj = imax(dumjj, 1);
    dumjj = imin(dumjj, 9);
    tab.rdumjj = rdumjj; tab.dumjj = dumjj;
  }

  // p3_table3_impl.hpp:70-83
  static Real apply_table(const Real (&table)[300][10], const Table3<Real>& tab3) {
    const Real rdumii_m_dumii = tab3.rdumii - Real(tab3.dumii);
    const Real t_im1_jm1 = table[tab3.dumii-1][tab3.dumjj-1];
    const Real dum1 = (t_im1_jm1 + rdumii_m_dumii * (table[tab3.dumii][tab3.dumjj-1] - t_im1_jm1));
    const Real t_im1_j = table[tab3.dumii-1][tab3.dumjj];
    const Real dum2 = (t_im1_j + rdumii_m_dumii * (table[tab3.dumii][tab3.dumjj] - t_im1_j));
    return dum1 + (tab3.rdumjj - Real(tab3.dumjj)) * (dum2 - dum1);
  }

  // p3_ice_collection_impl.hpp:12-52
  static void ice_cldliq_collection(Real rho, Real temp, Real rhofaci, Real table_val_qc2qi_collect,
      Real qi_incld, Real qc_incld, Real ni_incld, Real nc_incld,
      Real& qc2qi_collect_tend, Real& nc_collect_tend, Real& qc2qr_ice_shed_tend, Real& ncshdc,
      const Runtime& ro) {
    const Real qsmall = C::QSMALL, tmelt = C::Tmelt;
    const bool t_is_negative = temp <= tmelt;
    const bool both_ge_small = qi_incld >= qsmall && qc_incld >= qsmall;
    const bool both_ge_small_pos_t = both_ge_small && !t_is_negative;
    const Real f = ro.cldliq_to_ice_collection_factor;
    const Real inv_dropmass = C::ONE/C::dropmass;
    if (both_ge_small && t_is_negative)
      qc2qi_collect_tend = rhofaci * table_val_qc2qi_collect * qc_incld * f * rho * ni_incld;
    if (both_ge_small)
      nc_collect_tend = rhofaci * table_val_qc2qi_collect * nc_incld * f * rho * ni_incld;
    if (both_ge_small_pos_t) {
      qc2qr_ice_shed_tend = rhofaci * table_val_qc2qi_collect * qc_incld * f * rho * ni_incld;
      nc_collect_tend = rhofaci * table_val_qc2qi_collect * nc_incld * f * rho * ni_incld;
      ncshdc = qc2qr_ice_shed_tend*inv_dropmass;
    }
  }

  // p3_ice_collection_impl.hpp:57-103
  static void ice_rain_collection(Real rho, Real temp, Real rhofaci, Real logn0r,
      Real table_val_nr_collect, Real table_val_qr2qi_collect, Real qi_incld, Real ni_incld, Real qr_incld,
      Real& qr2qi_collect_tend, Real& nr_collect_tend, const Runtime& ro) {
    const Real qsmall = C::QSMALL, tmelt = C::Tmelt;
    const bool t_is_negative = temp <= tmelt;
    const bool both_ge_small = qi_incld >= qsmall && qr_incld >= qsmall;
    const Real ten = 10.0;
    const Real f = ro.rain_to_ice_collection_factor;
    if (both_ge_small && t_is_negative) {
      qr2qi_collect_tend = pm::pow(ten, table_val_qr2qi_collect + logn0r) * rho * rhofaci * f * ni_incld;
      nr_collect_tend    = pm::pow(ten, table_val_nr_collect + logn0r) * rho * rhofaci * f * ni_incld;
    }
    if (both_ge_small && !t_is_negative)
      nr_collect_tend    = pm::pow(ten, table_val_nr_collect + logn0r) * rho * rhofaci * f * ni_incld;
  }

  // p3_ice_collection_impl.hpp:107-153
  static void ice_self_collection(Real rho, Real rhofaci, Real table_val_ni_self_collect, Real eii,
      Real qm_incld, Real qi_incld, Real ni_incld, Real& ni_selfcollect_tend) {
    const Real qsmall = C::QSMALL;
    const bool qm_incld_positive = qm_incld > 0;
    const bool qi_incld_ge_small = qi_incld >= qsmall;
    Real tmp1 = 0, Eii_fact = 0;
    if (qi_incld_ge_small) {
      if (qm_incld_positive) tmp1 = qm_incld/qi_incld;
      if (tmp1 < sp(0.6) && qm_incld_positive) Eii_fact = 1;
      if (tmp1 >= sp(0.6) && tmp1 < sp(0.9) && qm_incld_positive) Eii_fact = 1 - (tmp1-sp(0.6))/sp(0.3);
      if (tmp1 >= sp(0.9) && qm_incld_positive) Eii_fact = 0;
      if (!qm_incld_positive) Eii_fact = 1;
      ni_selfcollect_tend = table_val_ni_self_collect*rho*eii*Eii_fact*rhofaci*ni_incld*ni_incld;
    }
  }

  // p3_ice_melting_impl.hpp:14-55
  static void ice_melting(Real rho, Real T_atm, Real pres, Real rhofaci,
      Real table_val_qi2qr_melting, Real table_val_qi2qr_vent_melt, Real dv, Real sc, Real mu, Real kap,
      Real qv, Real qi_incld, Real ni_incld, Real& qi2qr_melt_tend, Real& ni2nr_melt_tend, int* err) {
    const Real Pi = C::Pi, QSMALL = C::QSMALL, Tmelt = C::Tmelt, latvap = C::LatVap, latice = C::LatIce;
    const bool has_melt_qi = (qi_incld >= QSMALL) && (T_atm > Tmelt);
    if (has_melt_qi) {
      const Real qsat0 = qv_sat_dry(Tmelt, pres, false, err);
      qi2qr_melt_tend = ( (table_val_qi2qr_melting+table_val_qi2qr_vent_melt*pm::cbrt(sc)*std::sqrt(rhofaci*rho/mu))
                          *((T_atm-Tmelt)*kap-rho*latvap*dv*(qsat0-qv))
                          * 2 * Pi /latice)*ni_incld;
      qi2qr_melt_tend = max(qi2qr_melt_tend, 0);
      ni2nr_melt_tend = qi2qr_melt_tend*(ni_incld/qi_incld);
    }
  }

  // p3_ice_cldliq_wet_growth_impl.hpp:14-82
  static void ice_cldliq_wet_growth(Real rho, Real temp, Real pres, Real rhofaci,
      Real table_val_qi2qr_melting, Real table_val_qi2qr_vent_melt, Real dv, Real kap, Real mu, Real sc,
      Real qv, Real qc_incld, Real qi_incld, Real ni_incld, Real qr_incld,
      bool& log_wetgrowth, Real& qr2qi_collect_tend, Real& qc2qi_collect_tend, Real& qc_growth_rate,
      Real& nr_ice_shed_tend, Real& qc2qr_ice_shed_tend, int* err) {
    const Real qsmall = C::QSMALL, tmelt = C::Tmelt, twopi = C::Pi*2, zero = 0, one = 1;
    const Real cpw = C::CpLiq, latvap = C::LatVap, latice = C::LatIce;
    const bool any_if = qi_incld >= qsmall && (qc_incld + qr_incld) >= sp(1.0e-6) && temp < tmelt;
    const bool any_if_col = any_if && (qc2qi_collect_tend + qr2qi_collect_tend) >= sp(1.0e-10);
    if (any_if) {
      const Real qsat0 = qv_sat_dry(tmelt, pres, false, err);
      qc_growth_rate = ((table_val_qi2qr_melting+table_val_qi2qr_vent_melt*pm::cbrt(sc)*std::sqrt(rhofaci*rho/mu))*
                        twopi*(rho*latvap*dv*(qsat0-qv)-(temp-tmelt)*kap)/
                        (latice+cpw*(temp-tmelt)))*ni_incld;
      qc_growth_rate = max(qc_growth_rate, zero);
      const Real dum = max(zero, (qc2qi_collect_tend+qr2qi_collect_tend)-qc_growth_rate);
      const bool dum_ge_small = dum >= sp(1.0e-10);
      if (dum_ge_small) {
        nr_ice_shed_tend = nr_ice_shed_tend+dum*sp(1.923e+6);
        const Real dum1 = one/(qc2qi_collect_tend+qr2qi_collect_tend);
        if (any_if_col) {
          // all three right-hand sides use the pre-update qc2qi_collect_tend (l.70-76: the first
          // two .set() calls read qc2qi_collect_tend before it is overwritten by the second)
          const Real qc2qi_old = qc2qi_collect_tend;
          qc2qr_ice_shed_tend = qc2qr_ice_shed_tend+dum*qc2qi_old*dum1;
          qc2qi_collect_tend  = max(0, qc2qi_old-dum*qc2qi_old*dum1);
          qr2qi_collect_tend  = max(0, qr2qi_collect_tend-dum*qr2qi_collect_tend*dum1);
        }
      }
      log_wetgrowth = any_if && dum_ge_small;
    }
  }

  // p3_ice_relaxation_timescale_impl.hpp:12-43
  static void ice_relaxation_timescale(Real rho, Real temp, Real rhofaci,
      Real table_val_qi2qr_melting, Real table_val_qi2qr_vent_melt, Real dv, Real mu, Real sc,
      Real qi_incld, Real ni_incld, Real& epsi, Real& epsi_tot) {
    const Real qsmall = C::QSMALL, tmelt = C::Tmelt, pi = C::Pi;
    const bool any_if = qi_incld >= qsmall && temp < tmelt;
    if (any_if) {
      epsi = ((table_val_qi2qr_melting+table_val_qi2qr_vent_melt*pm::cbrt(sc)*std::sqrt(rhofaci*rho/mu))*
              sp(2.0)*pi*rho*dv)*ni_incld;
      epsi_tot = epsi_tot+epsi;
    } else {
      epsi = 0;
    }
  }

  // p3_calc_rime_density_impl.hpp:17-76
  static void calc_rime_density(Real T_atm, Real rhofaci, Real table_val_qi_fallspd, Real acn,
      Real lamc, Real mu_c, Real qc_incld, Real qc2qi_collect_tend, Real& vtrmi1, Real& rho_qm_cloud) {
    const Real qsmall = C::QSMALL, T_zerodegc = C::T_zerodegc, bcn = C::bcn;
    const bool m = (qc2qi_collect_tend >= qsmall) && (T_atm < T_zerodegc);
    if (m) {
      vtrmi1 = table_val_qi_fallspd * rhofaci;
      if (qc_incld >= qsmall) {
        const Real Vt_qc = acn * pm::tgamma(4+bcn+mu_c) / (pm::pow(lamc, bcn) * pm::tgamma(4+mu_c));
        const Real D_c = (4 + mu_c) / lamc;
        const Real V_impact = std::abs(vtrmi1-Vt_qc);
        const Real inv_Tc = 1/min(sp(-0.001), T_atm-T_zerodegc);
        const Real Ri = max(1, min(sp(-0.5e+6) * D_c * V_impact * inv_Tc, 12));
        if (Ri <= sp(8.0)) rho_qm_cloud = (sp(0.051) + sp(0.114)*Ri - sp(0.0055)*square(Ri))*1000;
        else               rho_qm_cloud = sp(611.) + sp(72.25) * (Ri-8);
      } else {
        rho_qm_cloud = 400;
      }
    } else {
      vtrmi1 = 0;
      rho_qm_cloud = 400;
    }
  }

  // p3_cldliq_imm_freezing_impl.hpp:18-55 (inv_qc_relvar unused: sgs_var_coef = 1)
  static void cldliq_immersion_freezing(Real T_atm, Real lamc, Real mu_c, Real cdist1, Real qc_incld,
      Real /*inv_qc_relvar*/, Real& qc2qi_hetero_freeze_tend, Real& nc2ni_immers_freeze_tend,
      const Runtime& ro) {
    const Real qsmall = C::QSMALL, T_rainfrz = C::T_rainfrz, T_zerodegc = C::T_zerodegc;
    const Real CONS5 = C::CONS5, CONS6 = C::CONS6;
    if ((qc_incld >= qsmall) && (T_atm <= T_rainfrz)) {
      const Real expAimmDt = pm::exp(ro.immersion_freezing_exponent * (T_zerodegc - T_atm));
      const Real inv_lamc3 = cube(1/lamc);
      const Real sgs_var_coef = 1;
      qc2qi_hetero_freeze_tend = sgs_var_coef * CONS6 * cdist1 * pm::tgamma(7+mu_c) * expAimmDt * square(inv_lamc3);
      nc2ni_immers_freeze_tend = CONS5 * cdist1 * pm::tgamma(sp(4.0)+mu_c) * expAimmDt * inv_lamc3;
    }
  }

  // p3_rain_imm_freezing_impl.hpp:17-49
  static void rain_immersion_freezing(Real T_atm, Real lamr, Real mu_r, Real cdistr, Real qr_incld,
      Real& qr2qi_immers_freeze_tend, Real& nr2ni_immers_freeze_tend, const Runtime& ro) {
    const Real qsmall = C::QSMALL, T_rainfrz = C::T_rainfrz, T_zerodegc = C::T_zerodegc;
    const Real CONS5 = C::CONS5, CONS6 = C::CONS6;
    if ((qr_incld >= qsmall) && (T_atm <= T_rainfrz)) {
      qr2qi_immers_freeze_tend = CONS6 *
          pm::exp(pm::log(cdistr) + pm::log(pm::tgamma(sp(7.) + mu_r)) - sp(6.) * pm::log(lamr)) *
          pm::exp(ro.immersion_freezing_exponent * (T_zerodegc - T_atm));
      nr2ni_immers_freeze_tend = CONS5 *
          pm::exp(pm::log(cdistr) + pm::log(pm::tgamma(sp(4.) + mu_r)) - sp(3.) * pm::log(lamr)) *
          pm::exp(ro.immersion_freezing_exponent * (T_zerodegc - T_atm));
    }
  }

  // p3_calc_liq_relaxation_timescale_impl.hpp:17-47
  static void calc_liq_relaxation_timescale(const Tables& t, Real rho, Real f1r, Real f2r,
      Real dv, Real mu, Real sc, Real mu_r, Real lamr, Real cdistr, Real cdist, Real qr_incld, Real qc_incld,
      Real& epsr, Real& epsc) {
    const Real qsmall = C::QSMALL, pi = C::Pi;
    epsr = 0;
    if (qr_incld >= qsmall) {
      Table3<Real> table;
      lookup(mu_r, lamr, table);
      epsr = 2 * pi * cdistr * rho * dv *
             (f1r * pm::tgamma(mu_r + 2)/lamr +
              f2r * std::sqrt(rho/mu) * pm::cbrt(sc) * apply_table(t.revap_table_vals, table));
    }
    if (qc_incld >= qsmall) epsc = 2 * pi * rho * dv * cdist;
    else                    epsc = 0;
  }

  // p3_evaporate_rain_impl.hpp:13-64
  static Real rain_evap_tscale_weight(Real dt_over_tau) { return -pm::expm1(-dt_over_tau)/dt_over_tau; }
  static Real rain_evap_equilib_tend(Real A_c, Real ab, Real tau_eff, Real tau_r) { return -A_c/ab*tau_eff/tau_r; }
  static Real rain_evap_instant_tend(Real ssat_r, Real ab, Real tau_r) { return -ssat_r/(ab*tau_r); }

  // p3_evaporate_rain_impl.hpp:67-205
  static void evaporate_rain(Real qr_incld, Real qc_incld, Real nr_incld, Real qi_incld,
      Real cld_frac_l, Real cld_frac_r, Real qv, Real qv_prev, Real qv_sat_l, Real qv_sat_i,
      Real ab, Real abi, Real epsr, Real epsi_tot, Real t_atm, Real t_atm_prev, Real dqsdt, Real dt,
      Real& qr2qv_evap_tend, Real& nr_evap_tend) {
    qr2qv_evap_tend = 0;
    nr_evap_tend = 0;
    const Real inv_dt = 1/dt;
    const Real QSMALL = C::QSMALL, Tmelt = C::Tmelt, inv_cp = 1/C::Cpair, latvap = C::LatVap, latice = C::LatIce;
    const Real ssat_r = qv - qv_sat_l;
    Real cld_frac;
    if ((qc_incld + qi_incld) < sp(1.e-6)) cld_frac = 0; else cld_frac = cld_frac_l;
    const bool is_rain_evap = qr_incld >= QSMALL && ssat_r < 0 && cld_frac_r > cld_frac;
    if (is_rain_evap) {
      const Real tau_r = 1/epsr;
      Real eps_eff, A_c;
      if (t_atm < Tmelt) {
        eps_eff = epsr + epsi_tot*(1 + (latvap+latice)*inv_cp*dqsdt)/abi;
        A_c = (qv - qv_prev)*inv_dt - dqsdt*(t_atm-t_atm_prev)*inv_dt
              - (qv_sat_l - qv_sat_i)*(1 + (latvap+latice)*inv_cp*dqsdt)/abi*epsi_tot;
      } else {
        eps_eff = epsr;
        A_c = (qv - qv_prev)*inv_dt - dqsdt*(t_atm-t_atm_prev)*inv_dt;
      }
      if (eps_eff < sp(1e-20)) eps_eff = sp(1e-20);
      const Real tau_eff = 1/eps_eff;
      const bool is_qr_tiny = qr_incld < sp(1e-12) && qv/qv_sat_l < sp(0.999);
      if (is_qr_tiny) {
        qr2qv_evap_tend = qr_incld*inv_dt;
      } else {
        const Real tscale_weight = rain_evap_tscale_weight(dt/tau_eff);
        const Real equilib_tend  = rain_evap_equilib_tend(A_c, ab, tau_eff, tau_r);
        const Real instant_tend  = rain_evap_instant_tend(ssat_r, ab, tau_r);
        qr2qv_evap_tend = instant_tend*tscale_weight + equilib_tend*(1-tscale_weight);
      }
      if (qr2qv_evap_tend > -ssat_r*inv_dt/ab) qr2qv_evap_tend = -ssat_r*inv_dt/ab;
      if (qr2qv_evap_tend < 0) qr2qv_evap_tend = 0;
      if (qr2qv_evap_tend > qr_incld*inv_dt) qr2qv_evap_tend = qr_incld*inv_dt;
      qr2qv_evap_tend = qr2qv_evap_tend*(cld_frac_r-cld_frac)/cld_frac_r;
      nr_evap_tend = qr2qv_evap_tend*(nr_incld/qr_incld);
    }
  }

  // p3_ice_deposition_sublimation_impl.hpp:13-67
  static void ice_deposition_sublimation(Real qi_incld, Real ni_incld, Real T_atm, Real qv_sat_l,
      Real qv_sat_i, Real epsi, Real abi, Real qv, Real inv_dt,
      Real& qv2qi_vapdep_tend, Real& qi2qv_sublim_tend, Real& ni_sublim_tend, Real& qc2qi_berg_tend) {
    const Real QSMALL = C::QSMALL, T_zerodegc = C::T_zerodegc;
    qc2qi_berg_tend = 0; qv2qi_vapdep_tend = 0; qi2qv_sublim_tend = 0; ni_sublim_tend = 0;
    if (qi_incld > QSMALL) {
      const Real qi_tend = min(epsi/abi, inv_dt) * (qv - qv_sat_i);
      const bool neg_qi_tend = qi_tend < 0;
      if (neg_qi_tend) {
        qi2qv_sublim_tend = -qi_tend;
        ni_sublim_tend = qi2qv_sublim_tend*(ni_incld/qi_incld);
      }
      const bool T_lt_frz = T_atm < T_zerodegc;
      if (T_lt_frz && !neg_qi_tend) qv2qi_vapdep_tend = qi_tend;
      if (T_lt_frz) qc2qi_berg_tend = max(epsi/abi*(qv_sat_l - qv_sat_i), 0);
    }
  }

  // p3_ice_nucleation_impl.hpp:12-63
  static void ice_nucleation(Real temp, Real inv_rho, Real ni, Real ni_activated, Real qv_supersat_i,
      Real inv_dt, bool do_predict_nc, bool do_prescribed_CCN,
      Real& qv2qi_nucleat_tend, Real& ni_nucleat_tend, const Runtime& ro) {
    const Real nsmall = C::NSMALL, tmelt = C::Tmelt, T_icenuc = C::Tmelt-sp(15.0), zero = 0;
    const Real piov3 = C::PIOV3;
    const Real mi0 = sp(4.0)*piov3*sp(900.0)*sp(1.e-18);
    const bool cond = temp < T_icenuc && qv_supersat_i >= sp(0.05);
    const bool do_log = (!do_predict_nc || do_prescribed_CCN);
    const bool any_if_log = cond && !do_log, any_if_not_log = cond && do_log;
    if (any_if_not_log) {
      Real dum = sp(0.005)*pm::exp(sp(ro.deposition_nucleation_exponent)*(tmelt-temp))*sp(1.0e3)*inv_rho;
      dum = min(dum, sp(1.0e5)*inv_rho);
      const Real N_nuc = max(zero, (dum-ni)*inv_dt);
      if (N_nuc >= nsmall) {
        const Real Q_nuc = max(zero, (dum-ni)*mi0*inv_dt);
        qv2qi_nucleat_tend = Q_nuc;
        ni_nucleat_tend = N_nuc;
      }
    } else if (any_if_log) {
      ni_nucleat_tend = max(zero, (ni_activated-ni)*inv_dt);
      qv2qi_nucleat_tend = ni_nucleat_tend*mi0;
    }
  }

  // p3_autoconversion_impl.hpp:13-60 (inv_qc_relvar unused: sgs_var_coef = 1)
  static void cloud_water_autoconversion(Real rho, Real qc_incld, Real nc_incld, Real /*inv_qc_relvar*/,
      Real& qc2qr_autoconv_tend, Real& nc2nr_autoconv_tend, Real& ncautr, const Runtime& ro) {
    const bool qc_not_small = qc_incld >= sp(1e-8);
    Real CONS3;
    if (ro.autoconversion_radius == sp(25.0e-6)) CONS3 = C::CONS3;
    else CONS3 = sp(1.0) / (C::CONS2 * pm::pow(ro.autoconversion_radius, sp(3.0)));
    if (qc_not_small) {
      const Real sgs_var_coef = 1;
      qc2qr_autoconv_tend = sgs_var_coef * ro.autoconversion_prefactor *
          pm::pow(qc_incld, ro.autoconversion_qc_exponent) *
          pm::pow(nc_incld * sp(1.e-6) * rho, -ro.autoconversion_nc_exponent);
      ncautr = qc2qr_autoconv_tend * CONS3;
      nc2nr_autoconv_tend = qc2qr_autoconv_tend * nc_incld / qc_incld;
    }
    if (qc2qr_autoconv_tend == 0) nc2nr_autoconv_tend = 0;
    if (nc2nr_autoconv_tend == 0) qc2qr_autoconv_tend = 0;
  }

  // p3_droplet_self_coll_impl.hpp:17-31
  static void droplet_self_collection(Real, Real, Real qc_incld, Real, Real, Real, Real& nc_selfcollect_tend) {
    if (qc_incld >= C::QSMALL) nc_selfcollect_tend = 0;
  }

  // p3_cloud_rain_acc_impl.hpp:18-57
  static void cloud_rain_accretion(Real /*rho*/, Real /*inv_rho*/, Real qc_incld, Real nc_incld, Real qr_incld,
      Real /*inv_qc_relvar*/, Real& qc2qr_accret_tend, Real& nc_accret_tend, const Runtime& ro) {
    const Real qsmall = C::QSMALL;
    const Real sgs_var_coef = 1;
    if ((qr_incld >= qsmall) && (qc_incld >= qsmall)) {
      if (ro.accretion_qc_exponent == ro.accretion_qr_exponent) {
        qc2qr_accret_tend = sgs_var_coef * ro.accretion_prefactor *
            pm::pow(qc_incld * qr_incld, ro.accretion_qr_exponent);
      } else {
        qc2qr_accret_tend = sgs_var_coef * ro.accretion_prefactor *
            pm::pow(qc_incld, ro.accretion_qc_exponent) * pm::pow(qr_incld, ro.accretion_qr_exponent);
      }
      nc_accret_tend = qc2qr_accret_tend * nc_incld / qc_incld;
      if (nc_accret_tend == 0) qc2qr_accret_tend = 0;
      if (qc2qr_accret_tend == 0) nc_accret_tend = 0;
    }
  }

  // p3_rain_self_collection_impl.hpp:12-58
  static void rain_self_collection(Real rho, Real qr_incld, Real nr_incld, Real& nr_selfcollect_tend,
      const Runtime& ro) {
    const Real qsmall = C::QSMALL, rho_h2o = C::RHO_H2O, pi = C::Pi;
    const Real dbrk = ro.rain_selfcollection_breakup_diameter;
    if (qr_incld >= qsmall) {
      const Real dum2 = pm::cbrt((qr_incld) / (pi * rho_h2o * nr_incld));
      Real dum;
      if (dum2 < dbrk) dum = 1;
      else             dum = 2 - pm::exp(2300 * (dum2 - dbrk));
      nr_selfcollect_tend = dum * ro.rain_selfcollection_prefactor * nr_incld * qr_incld * rho;
    }
  }

  // All per-level tendencies of part2 (p3_main_impl_part2.hpp:132-164), zero-initialised.
  struct Tend {
    Real qc2qr_accret_tend=0, qr2qv_evap_tend=0, qc2qr_autoconv_tend=0, nc_accret_tend=0,
         nc_selfcollect_tend=0, nc2nr_autoconv_tend=0, nr_selfcollect_tend=0, nr_evap_tend=0, ncautr=0,
         qi2qv_sublim_tend=0, nr_ice_shed_tend=0, qc2qi_hetero_freeze_tend=0, qr2qi_collect_tend=0,
         qc2qr_ice_shed_tend=0, qi2qr_melt_tend=0, qc2qi_collect_tend=0, qr2qi_immers_freeze_tend=0,
         qv2qi_nucleat_tend=0, ni2nr_melt_tend=0, nc_collect_tend=0, ncshdc=0, nc2ni_immers_freeze_tend=0,
         nr_collect_tend=0, ni_selfcollect_tend=0, ni_nucleat_tend=0, qv2qi_vapdep_tend=0,
         qc2qi_berg_tend=0, nr2ni_immers_freeze_tend=0, ni_sublim_tend=0, qc_growth_rate=0;
  };

  // p3_back_to_cell_average_impl.hpp:17-79
  static void back_to_cell_average(Real cld_frac_l, Real cld_frac_r, Real cld_frac_i, Tend& t) {
    const Real ir_cldm = min(cld_frac_i,cld_frac_r);
    const Real il_cldm = min(cld_frac_i,cld_frac_l);
    const Real lr_cldm = min(cld_frac_l,cld_frac_r);
    t.qc2qr_accret_tend *= lr_cldm;
    t.qr2qv_evap_tend *= cld_frac_r;
    t.qc2qr_autoconv_tend *= cld_frac_l;
    t.nc_accret_tend *= lr_cldm;
    t.nc_selfcollect_tend *= cld_frac_l;
    t.nc2nr_autoconv_tend *= cld_frac_l;
    t.nr_selfcollect_tend *= cld_frac_r;
    t.nr_evap_tend *= cld_frac_r;
    t.ncautr *= lr_cldm;
    t.qi2qv_sublim_tend *= cld_frac_i;
    t.nr_ice_shed_tend *= il_cldm;
    t.qc2qi_hetero_freeze_tend *= il_cldm;
    t.qr2qi_collect_tend *= ir_cldm;
    t.qc2qr_ice_shed_tend *= il_cldm;
    t.qi2qr_melt_tend *= cld_frac_i;
    t.qc2qi_collect_tend *= il_cldm;
    t.qr2qi_immers_freeze_tend *= cld_frac_r;
    t.ni2nr_melt_tend *= cld_frac_i;
    t.nc_collect_tend *= il_cldm;
    t.ncshdc *= il_cldm;
    t.nc2ni_immers_freeze_tend *= cld_frac_l;
    t.nr_collect_tend *= ir_cldm;
    t.ni_selfcollect_tend *= cld_frac_i;
    t.qv2qi_vapdep_tend *= cld_frac_i;
    t.nr2ni_immers_freeze_tend *= cld_frac_r;
    t.ni_sublim_tend *= cld_frac_i;
    t.qc2qi_berg_tend *= il_cldm;
  }

  // p3_conservation_impl.hpp:13-51
  static void cloud_water_conservation(Real qc, Real dt, Tend& t) {
    const Real sinks = (t.qc2qr_autoconv_tend+t.qc2qr_accret_tend+t.qc2qi_collect_tend+t.qc2qi_hetero_freeze_tend+t.qc2qr_ice_shed_tend+t.qc2qi_berg_tend)*dt;
    const Real sources = qc;
    Real ratio;
    const Real qtendsmall = C::QTENDSMALL;
    if (sinks > sources && sinks >= qtendsmall) {
      ratio = sources/sinks;
      t.qc2qr_autoconv_tend *= ratio;
      t.qc2qr_accret_tend *= ratio;
      t.qc2qi_collect_tend *= ratio;
      t.qc2qi_hetero_freeze_tend *= ratio;
      t.qc2qr_ice_shed_tend *= ratio;
      t.qc2qi_berg_tend *= ratio;
    } else {
      ratio = 1;
    }
    if (sources > qtendsmall) {
      t.qv2qi_vapdep_tend = t.qv2qi_vapdep_tend*(1-ratio);
      t.qi2qv_sublim_tend = t.qi2qv_sublim_tend*(1-ratio);
    }
  }

  // p3_conservation_impl.hpp:54-75
  static void rain_water_conservation(Real qr, Real dt, Tend& t) {
    const Real sinks   = (t.qr2qv_evap_tend+t.qr2qi_collect_tend+t.qr2qi_immers_freeze_tend)*dt;
    const Real sources = qr + (t.qc2qr_autoconv_tend+t.qc2qr_accret_tend+t.qi2qr_melt_tend+t.qc2qr_ice_shed_tend)*dt;
    if (sinks > sources && sinks >= C::QTENDSMALL) {
      const Real ratio = sources/sinks;
      t.qr2qv_evap_tend *= ratio;
      t.qr2qi_collect_tend *= ratio;
      t.qr2qi_immers_freeze_tend *= ratio;
    }
  }

  // p3_conservation_impl.hpp:78-99
  static void ice_water_conservation(Real qi, Real dt, Tend& t) {
    const Real sinks = (t.qi2qv_sublim_tend+t.qi2qr_melt_tend)*dt;
    const Real sources = qi + (t.qv2qi_vapdep_tend+t.qv2qi_nucleat_tend+t.qr2qi_collect_tend+t.qc2qi_collect_tend
                               + t.qr2qi_immers_freeze_tend+t.qc2qi_hetero_freeze_tend+t.qc2qi_berg_tend)*dt;
    if (sinks > sources && sinks >= C::QTENDSMALL) {
      const Real ratio = sources/sinks;
      t.qi2qv_sublim_tend *= ratio;
      t.qi2qr_melt_tend *= ratio;
    }
  }

  // p3_nc_conservation_impl.hpp:16-33
  static void nc_conservation(Real nc, Real dt, Tend& t) {
    const Real sink_nc = (t.nc_collect_tend + t.nc2ni_immers_freeze_tend + t.nc_accret_tend + t.nc2nr_autoconv_tend)*dt;
    const Real source_nc = nc + t.nc_selfcollect_tend*dt;
    if (sink_nc > source_nc) {
      const Real ratio = source_nc/sink_nc;
      t.nc_collect_tend *= ratio;
      t.nc2ni_immers_freeze_tend *= ratio;
      t.nc_accret_tend *= ratio;
      t.nc2nr_autoconv_tend *= ratio;
    }
  }

  // p3_nr_conservation_impl.hpp:16-33
  static void nr_conservation(Real nr, Real dt, Real nmltratio, Tend& t) {
    const Real sink_nr = (t.nr_collect_tend + t.nr2ni_immers_freeze_tend + t.nr_selfcollect_tend + t.nr_evap_tend)*dt;
    const Real source_nr = nr + (t.ni2nr_melt_tend*nmltratio + t.nr_ice_shed_tend + t.ncshdc + t.nc2nr_autoconv_tend)*dt;
    if (sink_nr > source_nr) {
      const Real ratio = source_nr/sink_nr;
      t.nr_collect_tend *= ratio;
      t.nr2ni_immers_freeze_tend *= ratio;
      t.nr_selfcollect_tend *= ratio;
      t.nr_evap_tend *= ratio;
    }
  }

  // p3_ni_conservation_impl.hpp:16-32
  static void ni_conservation(Real ni, Real dt, Tend& t) {
    const Real sink_ni = (t.ni2nr_melt_tend + t.ni_sublim_tend + t.ni_selfcollect_tend)*dt;
    const Real source_ni = ni + (t.ni_nucleat_tend+t.nr2ni_immers_freeze_tend+t.nc2ni_immers_freeze_tend)*dt;
    if (sink_ni > source_ni) {
      const Real ratio = source_ni/sink_ni;
      t.ni2nr_melt_tend *= ratio;
      t.ni_sublim_tend *= ratio;
      t.ni_selfcollect_tend *= ratio;
    }
  }

  // p3_ice_supersat_conservation_impl.hpp:16-45
  static void ice_supersat_conservation(Real& qv2qi_vapdep_tend, Real& qv2qi_nucleat_tend, Real cld_frac_i,
      Real qv, Real qv_sat_i, Real t_atm, Real dt, Real qi2qv_sublim_tend, Real qr2qv_evap_tend) {
    const Real qsmall = C::QSMALL, cp = C::CP, rv = C::RH2O, latvap = C::LatVap, latice = C::LatIce;
    const Real latsublim2 = (latvap+latice)*(latvap+latice);
    const Real qv_sink = qv2qi_vapdep_tend + qv2qi_nucleat_tend;
    if (qv_sink > qsmall && cld_frac_i > sp(1e-20)) {
      Real qv_avail = (qv + (qi2qv_sublim_tend+qr2qv_evap_tend)*dt - qv_sat_i) /
                      (1 + latsublim2*qv_sat_i / (cp*rv*square(t_atm)) ) / dt;
      qv_avail = max(qv_avail, 0);
      if (qv_sink > qv_avail) {
        const Real fract = qv_avail / qv_sink;
        qv2qi_nucleat_tend = qv2qi_nucleat_tend * fract;
        qv2qi_vapdep_tend = qv2qi_vapdep_tend * fract;
      }
    }
  }

  // p3_prevent_liq_supersaturation_impl.hpp:17-80
  static void prevent_liq_supersaturation(Real pres, Real t_atm, Real qv, Real dt,
      Real qv2qi_vapdep_tend, Real qinuc, Real& qi2qv_sublim_tend, Real& qr2qv_evap_tend, int* err) {
    const Real inv_cp = C::INV_CP, rv = C::RV, qsmall = C::QSMALL, latvap = C::LatVap, latice = C::LatIce;
    const Real qv_sources = qi2qv_sublim_tend + qr2qv_evap_tend;
    if (!(qv_sources >= qsmall)) return;
    const Real qv_sinks = qv2qi_vapdep_tend + qinuc;
    const Real qv_endstep = qv - qv_sinks*dt + qv_sources*dt;
    (void)qv_endstep;
    const Real T_endstep = t_atm + ( (qv_sinks-qi2qv_sublim_tend)*(latvap+latice)*inv_cp
                                     - qr2qv_evap_tend*latvap*inv_cp )*dt;
    const Real qsl = qv_sat_dry(T_endstep, pres, false, err);
    const Real A = latvap*qsl*dt*inv_cp/(rv*T_endstep*T_endstep)
                   * ((latvap+latice)*qi2qv_sublim_tend + latvap*qr2qv_evap_tend);
    Real frac = (qsl-qv+qv_sinks*dt + A)/(qv_sources*dt + A);
    frac = max(0, frac);
    frac = min(1, frac);
    qi2qv_sublim_tend = frac*qi2qv_sublim_tend;
    qr2qv_evap_tend = frac*qr2qv_evap_tend;
  }

  // p3_update_prognostics_impl.hpp:12-86
  static void update_prognostic_ice(const Tend& t, Real inv_exner, bool do_predict_nc, bool log_wetgrowth,
      Real dt, Real nmltratio, Real rho_qm_cloud,
      Real& th_atm, Real& qv, Real& qi, Real& ni, Real& qm, Real& bm, Real& qc, Real& nc, Real& qr, Real& nr) {
    const Real QSMALL = C::QSMALL, INV_RHO_RIMEMAX = C::INV_RHO_RIMEMAX, latvap = C::LatVap, latice = C::LatIce;
    qc = qc + (-t.qc2qi_hetero_freeze_tend-t.qc2qi_collect_tend-t.qc2qr_ice_shed_tend-t.qc2qi_berg_tend)*dt;
    if (do_predict_nc) nc = nc + (-t.nc_collect_tend-t.nc2ni_immers_freeze_tend)*dt;
    qr = qr + (-t.qr2qi_collect_tend+t.qi2qr_melt_tend-t.qr2qi_immers_freeze_tend+t.qc2qr_ice_shed_tend)*dt;
    nr = nr + (-t.nr_collect_tend-t.nr2ni_immers_freeze_tend+nmltratio*t.ni2nr_melt_tend+t.nr_ice_shed_tend+t.ncshdc)*dt;
    if (qi >= QSMALL) {
      bm = bm - ((t.qi2qv_sublim_tend + t.qi2qr_melt_tend) / qi) * dt * bm;
      qm = qm - ((t.qi2qv_sublim_tend + t.qi2qr_melt_tend) * qm / qi) * dt;
      qi = qi - (t.qi2qv_sublim_tend + t.qi2qr_melt_tend) * dt;
    }
    const Real dum = (t.qr2qi_collect_tend + t.qc2qi_collect_tend + t.qr2qi_immers_freeze_tend + t.qc2qi_hetero_freeze_tend) * dt;
    qi = qi + (t.qv2qi_vapdep_tend + t.qv2qi_nucleat_tend + t.qc2qi_berg_tend) * dt + dum;
    qm = qm + dum;
    bm = bm + (t.qr2qi_collect_tend * INV_RHO_RIMEMAX + t.qc2qi_collect_tend / rho_qm_cloud +
               (t.qr2qi_immers_freeze_tend + t.qc2qi_hetero_freeze_tend) * INV_RHO_RIMEMAX) * dt;
    ni = ni + (t.ni_nucleat_tend - t.ni2nr_melt_tend - t.ni_sublim_tend - t.ni_selfcollect_tend +
               t.nr2ni_immers_freeze_tend + t.nc2ni_immers_freeze_tend) * dt;
    if (qm < 0) { qm = 0; bm = 0; }
    if (log_wetgrowth) { qm = qi; bm = qm * INV_RHO_RIMEMAX; }
    qv = qv + (-t.qv2qi_vapdep_tend+t.qi2qv_sublim_tend-t.qv2qi_nucleat_tend)*dt;
    const Real INV_CP = C::INV_CP;
    th_atm = th_atm + inv_exner * ((t.qv2qi_vapdep_tend - t.qi2qv_sublim_tend + t.qv2qi_nucleat_tend) * (latvap+latice) * INV_CP +
                                   (t.qr2qi_collect_tend + t.qc2qi_collect_tend + t.qc2qi_hetero_freeze_tend + t.qr2qi_immers_freeze_tend -
                                    t.qi2qr_melt_tend + t.qc2qi_berg_tend) * latice * INV_CP) * dt;
  }

  // p3_update_prognostics_impl.hpp:89-124
  static void update_prognostic_liquid(const Tend& t, bool do_predict_nc, bool do_prescribed_CCN,
      Real inv_rho, Real inv_exner, Real dt, Real& th_atm, Real& qv, Real& qc, Real& nc, Real& qr, Real& nr) {
    const Real NCCNST = C::NCCNST, INV_CP = C::INV_CP, latvap = C::LatVap;
    qc = qc + (-t.qc2qr_accret_tend-t.qc2qr_autoconv_tend)*dt;
    qr = qr + (t.qc2qr_accret_tend+t.qc2qr_autoconv_tend-t.qr2qv_evap_tend)*dt;
    if (do_predict_nc || do_prescribed_CCN) nc = nc + (-t.nc_accret_tend-t.nc2nr_autoconv_tend+t.nc_selfcollect_tend)*dt;
    else nc = NCCNST * inv_rho;
    if (C::IPARAM == 1 || C::IPARAM == 2) nr = nr + (sp(0.5) * t.nc2nr_autoconv_tend - t.nr_selfcollect_tend - t.nr_evap_tend) * dt;
    else nr = nr + (t.ncautr - t.nr_selfcollect_tend - t.nr_evap_tend) * dt;
    qv = qv + t.qr2qv_evap_tend *dt;
    th_atm = th_atm + inv_exner*(-t.qr2qv_evap_tend * latvap * INV_CP) * dt;
  }
};  // struct Procs

} // namespace p3o
#endif
