// p3_column.cuh — one P3 column on one warp (B200 layout: levels strided across the 32 lanes).
//
// Replaces the body of the "p3 main loop" team lambda (p3/impl/p3_main_impl.hpp:108-325): where the
// reference runs 8 team-parallel phases separated by 15 team_barrier sites and streams 44 per-column
// temporaries through a global-memory WorkspaceManager, this driver makes three passes over a column
// whose working set lives in shared memory / registers:
//
//   pass 1  p3_main_init + p3_main_part1 + p3_main_part2 fused per level (all part1->part2
//           temporaries stay in registers), prognostics parked in shared memory;
//   sedim.  cloud / rain / ice CFL sub-stepping on the shared-memory copy: fall speeds per lane,
//           Co_max by warp reduction, upwind flux from the level above through a shared flux row;
//   pass 3  homogeneous_freezing + p3_main_part3 fused per level, single write-back to HBM.
//
// Why fusing part1 and part2 is exact: the reference skips part2 for a column iff no level can
// nucleate and no level holds hydrometeors after part1 (p3_main_impl.hpp:237).  For such a column
// every non-skip_all level has zero in-cloud mass and T >= 273.15 K, so all 29 tendencies are exactly
// zero and part2 leaves every prognostic bit-identical; the only visible difference would be the
// three history arrays, whose stores are therefore deferred until the column flags are known.
//
// WP is the warp policy: WarpDev (32 lanes, shuffles) on the GPU, WarpSim (1 lane) in the
// host-side debug build of this same source (tests/hostsim; never part of the product library).
#ifndef P3B200_COLUMN_CUH
#define P3B200_COLUMN_CUH

#include "p3_physics.cuh"

namespace p3b200 {

enum FieldId {
  F_QC = 0, F_NC, F_QR, F_NR, F_QI, F_QM, F_NI, F_BM, F_QV, F_TH,
  F_NC_NUCEAT_TEND, F_NCCN, F_NI_ACTIVATED, F_INV_QC_RELVAR, F_CLD_FRAC_I, F_CLD_FRAC_L, F_CLD_FRAC_R,
  F_PRES, F_DZ, F_DPRES, F_INV_EXNER, F_QV_PREV, F_T_PREV,
  F_QV2QI_DEPOS_TEND, F_DIAG_EFF_RADIUS_QC, F_DIAG_EFF_RADIUS_QI, F_DIAG_EFF_RADIUS_QR, F_RHO_QI,
  F_PRECIP_LIQ_FLUX, F_PRECIP_ICE_FLUX, F_PRECIP_TOTAL_TEND, F_NEVAPR,
  F_LIQ_ICE_EXCHANGE, F_VAP_LIQ_EXCHANGE, F_VAP_ICE_EXCHANGE,
  F_PRECIP_LIQ_SURF, F_PRECIP_ICE_SURF, F_NUM
};

enum DiagId {
  D_NUC_POSSIBLE = 0, D_HYDRO_PART1, D_HYDRO_PART2,
  D_QC_KTOP, D_QC_KBOT, D_QC_NSUB, D_QR_KTOP, D_QR_KBOT, D_QR_NSUB, D_QI_KTOP, D_QI_KBOT, D_QI_NSUB,
  D_ERROR, D_NUM = 16
};

constexpr int MAX_NLEV = 256;
constexpr int NUM_SMEM_ROWS = 16;   // rows of nk Reals per column, see ColumnSmem

// Kernel arguments (by value in param space).
template <typename Real>
struct MainArgs {
  Real*   f[F_NUM];
  int64_t stride[F_NUM];
  int     ncol, nk;
  Real    dt;
  int     predictNc, prescribedCCN, zero_flux_top;
  Tables<Real>  tb;
  Runtime<Real> ro;
  int*    diag;            // [ncol][D_NUM]
  unsigned int* queue;     // work-queue counter (persistent warps)
};

// Shared-memory working set of one column: NUM_SMEM_ROWS rows of `nk` Reals.
template <typename Real>
struct ColumnSmem {
  Real *qc, *nc, *qr, *nr, *qi, *qm, *ni, *bm, *qv, *th;   // prognostic state
  Real *rho, *inv_dz;
  Real *aux;     // pass1: T_atm at init (for acn);  cloud sed: -acn once computed;  rain/ice sed: rhofac{r,i} (0 = not yet)
  Real *Vq, *Vn, *flux;   // sedimentation rows; during pass 1 they park the three history outputs
  P3_HD void bind(Real* base, int nk) {
    Real** rows[NUM_SMEM_ROWS] = {&qc, &nc, &qr, &nr, &qi, &qm, &ni, &bm, &qv, &th, &rho, &inv_dz, &aux, &Vq, &Vn, &flux};
    for (int r = 0; r < NUM_SMEM_ROWS; ++r) *rows[r] = base + r * nk;
  }
};

// ---- lane-private bit sets indexed by level slot (k / WP::size) -----------------------------------
template <int NWORDS>
struct LaneBits {
  unsigned w[NWORDS];
  P3_HD void clear() { for (int i = 0; i < NWORDS; ++i) w[i] = 0u; }
  P3_HD void set(int j) { w[j >> 5] |= (1u << (j & 31)); }
  P3_HD bool get(int j) const { return (w[j >> 5] >> (j & 31)) & 1u; }
};

template <typename Real, typename WP, int JM_>
struct Column {
  static constexpr int JM = JM_;   // level slots per lane held in registers by sedimentation: JM*WP::size >= nk
  using P = Phys<Real>;
  using C = Cst<Real>;
  static constexpr int WS = WP::size;
  static constexpr int NSLOT = (MAX_NLEV + WS - 1) / WS;
  using Bits = LaneBits<(NSLOT + 31) / 32>;
  static P3_HD Real sp(double x) { return static_cast<Real>(x); }

  // ---------------------------------------------------------------------------------------------
  // find_top / find_bottom (p3_find_impl.hpp:57-92, 17-52) for kdir = -1 over the whole column.
  static P3_HD bool find_range(const WP& w, int lane, int nk, const Real* q, int& k_top, int& k_bot) {
    int lo = MAX_NLEV + 1, hi = -1;
    for (int k = lane; k < nk; k += WS)
      if (q[k] >= C::QSMALL) { if (k < lo) lo = k; hi = k; }
    k_top = w.min_int(lo);
    k_bot = w.max_int(hi);
    return k_bot >= 0;
  }

  enum { SP_CLOUD = 0, SP_RAIN = 1, SP_ICE = 2 };

  // One first-order upwind sub-step of one field held in the register window
  // (calc_first_order_upwind_step, p3_upwind_impl.hpp:18-91, kdir = -1: inflow from level k-1).
  // Slot j of a lane holds level kw + lane + j*WS, so the level above lives in the previous lane of the
  // same slot (one shuffle) or, for lane 0, in the last lane of the previous slot.  flux[] returns the
  // fluxes of this field (needed for the surface accumulation and precip_liq_flux).
  static P3_HD void upwind_window(const WP& w, int lane, int J, int kw, int k_temp, Real dt_sub,
                                  const Real (&V)[JM], Real (&r)[JM], const Real (&rho)[JM],
                                  const Real (&inv_rho)[JM], const Real (&inv_dz)[JM], Real (&flux)[JM]) {
    Real prev_last = 0;
#pragma unroll
    for (int j = 0; j < JM; ++j) {
      if (j < J) {
        const int k = kw + lane + j * WS;
        const bool in = k <= k_temp;
        const Real f = in ? V[j] * r[j] * rho[j] : Real(0);
        Real up = w.shfl_up1(f);
        if (lane == 0) up = prev_last;
        prev_last = w.shfl_from(f, WS - 1);
        if (in) {
          const Real fluxdiv = (up - f) * inv_dz[j];
          r[j] += fluxdiv * dt_sub * inv_rho[j];
        }
        flux[j] = f;
      }
    }
  }

  // cloud_sedimentation (p3_cloud_sed_impl.hpp:17-147), rain_sedimentation (p3_rain_sed_impl.hpp:39-170),
  // ice_sedimentation (p3_ice_sed_impl.hpp:69-205) share this CFL sub-step loop; SP selects the
  // fall-speed physics.
  //
  // Layout: the levels [k_qxtop, nk) that the hydrometeor can reach form a WINDOW anchored at k_qxtop;
  // lane l keeps levels k_qxtop + l + j*WS (j = 0..J-1) entirely in registers for the whole sub-step
  // loop (state, rho, 1/rho, 1/dz, cloud fraction, lazily evaluated density factor), slots being
  // activated as the falling front (k_qxbot, +1 level per sub-step) reaches them.  A sub-step is then:
  // fall speeds per lane, one warp max for the Courant number, and per field one shuffle for the
  // inflow flux — no shared-memory round trips and no divisions that the reference does not have.
  //
  // In-cloud values are re-derived from the cell values: on the first sub-step they equal what
  // calculate_incloud_mixingratios left at the end of part1/part2 (limits included), afterwards
  // q/cld_frac (the reference's end-of-sub-step refresh).
  template <int SP>
  static P3_HD void sediment(const WP& w, int lane, const MainArgs<Real>& a, int64_t col,
                             const ColumnSmem<Real>& s, Real inv_dt, Real& surf_accum_out, int* dloc, int& err) {
    const int nk = a.nk, kbot = nk - 1;
    const Real qsmall = C::QSMALL;
    Real* sq = (SP == SP_CLOUD) ? s.qc : (SP == SP_RAIN) ? s.qr : s.qi;
    Real* sn = (SP == SP_CLOUD) ? s.nc : (SP == SP_RAIN) ? s.nr : s.ni;
    const int cf_id = (SP == SP_CLOUD) ? F_CLD_FRAC_L : (SP == SP_RAIN) ? F_CLD_FRAC_R : F_CLD_FRAC_I;
    const Real* cf_p = a.f[cf_id] + col * a.stride[cf_id];
    const Real* cfl_p = a.f[F_CLD_FRAC_L] + col * a.stride[F_CLD_FRAC_L];
    Real* plf = a.f[F_PRECIP_LIQ_FLUX] + col * a.stride[F_PRECIP_LIQ_FLUX];
    int k_qxtop, k_qxbot;
    const bool present = find_range(w, lane, nk, sq, k_qxtop, k_qxbot);
    const int dbase = (SP == SP_CLOUD) ? D_QC_KTOP : (SP == SP_RAIN) ? D_QR_KTOP : D_QI_KTOP;
    int nsub = 0;
    surf_accum_out = 0;
    if (!present) { dloc[dbase] = -1; dloc[dbase + 1] = -1; dloc[dbase + 2] = 0; return; }

    const int kw = k_qxtop;
    Real q[JM], n[JM], qm[JM], bm[JM], rho[JM], inv_rho[JM], inv_dz[JM], cf[JM], fac[JM], acc[JM];
    Real Vq[JM], Vn[JM], flux[JM], ftmp[JM];
    int J = 0;                   // slots activated so far (warp-uniform)
    Real dt_left = a.dt, prt_local = 0;
    bool first = true;
    while (dt_left > C::dt_left_tol) {
      const int k_temp = (k_qxbot == kbot) ? k_qxbot : k_qxbot + 1;
      // activate the slots the front reaches in this sub-step
      const int Jn = (k_temp - kw) / WS + 1;
#pragma unroll
      for (int j = 0; j < JM; ++j) {
        if (j >= J && j < Jn) {
          const int k = kw + lane + j * WS;
          if (k < nk) {
            q[j] = sq[k]; n[j] = sn[k];
            if (SP == SP_ICE) { qm[j] = s.qm[k]; bm[j] = s.bm[k]; }
            rho[j] = s.rho[k]; inv_rho[j] = 1 / rho[j]; inv_dz[j] = s.inv_dz[k];
            cf[j] = P3_LDG(cf_p + k);
            if (SP == SP_RAIN) acc[j] = plf[k + 1];
          } else {
            q[j] = 0; n[j] = 0; rho[j] = 1; inv_rho[j] = 1; inv_dz[j] = 0; cf[j] = 1;
            if (SP == SP_ICE) { qm[j] = 0; bm[j] = 0; }
            if (SP == SP_RAIN) acc[j] = 0;
          }
          fac[j] = 0;
        }
      }
      if (Jn > J) J = Jn;

      Real Co_loc = 0;
#pragma unroll
      for (int j = 0; j < JM; ++j) {
        if (j < J) {
          const int k = kw + lane + j * WS;
          Real vq = 0, vn = 0;
          if (k <= k_qxbot) {
            if (SP == SP_CLOUD) {
              Real qc_i, nc_i;
              if (first) P::incloud_liq(q[j], n[j], 1 / cf[j], qc_i, nc_i);
              else { qc_i = q[j] / cf[j]; nc_i = n[j] / cf[j]; }
              if (qc_i > qsmall) {
                Real mu_c, lamc, d0, d1;
                P::cloud_dsd2(qc_i, nc_i, rho[j], mu_c, lamc, d0, d1, false);
                n[j] = nc_i * cf[j];
                if (fac[j] == 0) {   // acn = g*rhow/(18*mu_air(T_atm at init)), p3_main_impl_part1.hpp:108-109
                  const Real T0 = s.aux[k];
                  const Real dum = sp(1.496e-6) * std::pow(T0, sp(1.5)) / (T0 + 120);
                  fac[j] = C::gravit * C::RHO_H2O / (18 * dum);
                }
                const Real acn = fac[j];
                const Real bcn = C::bcn;
                const Real dum = 1 / std::pow(lamc, bcn);
                vq = acn * std::tgamma(4 + bcn + mu_c) * dum / std::tgamma(mu_c + 4);
                if (a.predictNc) vn = acn * std::tgamma(1 + bcn + mu_c) * dum / std::tgamma(mu_c + 1);
                const Real Co = vq * dt_left * inv_dz[j];
                if (Co > Co_loc) Co_loc = Co;
              }
            } else if (SP == SP_RAIN) {
              Real qr_i, nr_i;
              if (first) P::incloud_rain(q[j], n[j], 1 / cf[j], qr_i, nr_i);
              else { qr_i = q[j] / cf[j]; nr_i = n[j] / cf[j]; }
              if (qr_i > qsmall) {
                // compute_rain_fall_velocity: p3_rain_sed_impl.hpp:17-34
                Real mu_r = 0, lamr;
                P::rain_dsd2(qr_i, nr_i, mu_r, lamr, a.ro);
                typename P::Loc3 t3;
                P::lookup3(mu_r, lamr, t3);
                if (fac[j] == 0) fac[j] = std::pow(C::RHO_1000MB * inv_rho[j], sp(.54));   // rhofacr
                vq = P::apply_table3(a.tb.vm, t3) * fac[j];
                vn = P::apply_table3(a.tb.vn, t3) * fac[j];
                n[j] = nr_i * cf[j];
                const Real Co = vq * dt_left * inv_dz[j];
                if (Co > Co_loc) Co_loc = Co;
              }
            } else {
              Real qi_i, ni_i, qm_i, bm_i;
              if (first) P::incloud_ice(q[j], n[j], qm[j], bm[j], 1 / cf[j], 1 / P3_LDG(cfl_p + k), qi_i, ni_i, qm_i, bm_i);
              else { qi_i = q[j] / cf[j]; ni_i = n[j] / cf[j]; qm_i = qm[j] / cf[j]; bm_i = bm[j] / cf[j]; }
              if (qi_i > qsmall) {
                ni_i = P::rmax(ni_i, C::NSMALL);
                const Real rhop = P::calc_bulk_rho_rime(qi_i, qm_i, bm_i, a.ro);
                qm[j] = qm_i * cf[j];
                bm[j] = bm_i * cf[j];
                typename P::IceLoc ti;
                P::lookup_ice(qi_i, ni_i, qm_i, rhop, ti);
                const Real tv_ni_fallspd = P::apply_table_ice(0, a.tb.ice, ti);
                const Real tv_qi_fallspd = P::apply_table_ice(1, a.tb.ice, ti);
                const Real tv_ni_lammax  = P::apply_table_ice(6, a.tb.ice, ti);
                const Real tv_ni_lammin  = P::apply_table_ice(7, a.tb.ice, ti);
                ni_i = P::rmin(ni_i, tv_ni_lammax * ni_i);
                ni_i = P::rmax(ni_i, tv_ni_lammin * ni_i);
                n[j] = ni_i * cf[j];
                if (fac[j] == 0) fac[j] = std::pow(C::RHO_600MB * inv_rho[j], sp(.54));     // rhofaci
                vq = a.ro.ice_sedimentation_factor * tv_qi_fallspd * fac[j];
                vn = a.ro.ice_sedimentation_factor * tv_ni_fallspd * fac[j];
                const Real Co = vq * dt_left * inv_dz[j];
                if (Co > Co_loc) Co_loc = Co;
              }
            }
          }
          Vq[j] = vq; Vn[j] = vn;
        }
      }
      const Real Co_max = w.max_real(Co_loc);
      // generalized_sedimentation: p3_upwind_impl.hpp:97-129
      if (!(Co_max >= 0)) err |= 4;
      const int tmpint1 = static_cast<int>(Co_max + 1);
      const Real dt_sub = dt_left / tmpint1;
      if (SP == SP_ICE) {
        upwind_window(w, lane, J, kw, k_temp, dt_sub, Vq, bm, rho, inv_rho, inv_dz, ftmp);
        upwind_window(w, lane, J, kw, k_temp, dt_sub, Vq, qm, rho, inv_rho, inv_dz, ftmp);
      }
      if (SP != SP_CLOUD || a.predictNc) upwind_window(w, lane, J, kw, k_temp, dt_sub, Vn, n, rho, inv_rho, inv_dz, ftmp);
      upwind_window(w, lane, J, kw, k_temp, dt_sub, Vq, q, rho, inv_rho, inv_dz, flux);
#pragma unroll
      for (int j = 0; j < JM; ++j) {
        if (j < J) {
          const int k = kw + lane + j * WS;
          if (k == kbot && k_qxbot == kbot) prt_local += flux[j] * dt_sub;
          // precip_liq_flux(k+1) += flux_qx(k) for k in [k_qxtop, new k_qxbot], p3_rain_sed_impl.hpp:139-153
          if (SP == SP_RAIN && k <= k_temp) acc[j] = acc[j] + flux[j];
        }
      }
      if (k_qxbot != kbot) k_qxbot += 1;
      dt_left -= dt_sub;
      first = false;
      ++nsub;
    }
    // write the window back
#pragma unroll
    for (int j = 0; j < JM; ++j) {
      if (j < J) {
        const int k = kw + lane + j * WS;
        if (k < nk) {
          sq[k] = q[j]; sn[k] = n[j];
          if (SP == SP_ICE) { s.qm[k] = qm[j]; s.bm[k] = bm[j]; }
          if (SP == SP_RAIN) plf[k + 1] = acc[j];
        }
      }
    }
    w.sync();
    const Real prt_accum = w.shfl_from(prt_local, (kbot - kw) % WS);   // only the lane owning kbot accumulated
    surf_accum_out = prt_accum * C::INV_RHO_H2O * inv_dt;
    dloc[dbase] = k_qxtop; dloc[dbase + 1] = k_qxbot; dloc[dbase + 2] = nsub;
  }

  // ---------------------------------------------------------------------------------------------
  static P3_HD void run(const WP& w, const MainArgs<Real>& a, int64_t col, Real* smem_base) {
    const int lane = w.lane();
    const int nk = a.nk;
    const Real qsmall = C::QSMALL, Tz = C::Tmelt, latvap = C::LatVap, latice = C::LatIce, inv_cp = C::INV_CP;
    const Real dt = a.dt, inv_dt = 1 / a.dt;
    const bool predictNc = a.predictNc != 0, prescribedCCN = a.prescribedCCN != 0;
    ColumnSmem<Real> s;
    s.bind(smem_base, nk);

    auto colp = [&](int f) -> Real* { return a.f[f] + col * a.stride[f]; };
    Real* const g_qc = colp(F_QC); Real* const g_nc = colp(F_NC); Real* const g_qr = colp(F_QR);
    Real* const g_nr = colp(F_NR); Real* const g_qi = colp(F_QI); Real* const g_qm = colp(F_QM);
    Real* const g_ni = colp(F_NI); Real* const g_bm = colp(F_BM); Real* const g_qv = colp(F_QV);
    Real* const g_th = colp(F_TH);
    const Real* const g_cfi = colp(F_CLD_FRAC_I); const Real* const g_cfl = colp(F_CLD_FRAC_L);
    const Real* const g_cfr = colp(F_CLD_FRAC_R); const Real* const g_pres = colp(F_PRES);
    const Real* const g_dz = colp(F_DZ); const Real* const g_dpres = colp(F_DPRES);
    const Real* const g_inv_exner = colp(F_INV_EXNER);
    const Real* const g_qv_prev = colp(F_QV_PREV); const Real* const g_t_prev = colp(F_T_PREV);
    const Real* const g_nccn = a.f[F_NCCN] ? colp(F_NCCN) : nullptr;
    const Real* const g_nc_nuceat = a.f[F_NC_NUCEAT_TEND] ? colp(F_NC_NUCEAT_TEND) : nullptr;
    const Real* const g_ni_act = a.f[F_NI_ACTIVATED] ? colp(F_NI_ACTIVATED) : nullptr;
    Real* const g_plf = colp(F_PRECIP_LIQ_FLUX); Real* const g_pif = colp(F_PRECIP_ICE_FLUX);

    int err = 0;
    bool nuc = false, hyd1 = false, hyd2 = false;
    Bits frz, p2;      // per level slot: T_atm < T_homogfrz (after part1); level ran part2
    frz.clear(); p2.clear();

    // ================= pass 1: init + part1 + part2, one level per lane at a time =================
    int slot = 0;
    for (int k = lane; k < nk; k += WS, ++slot) {
      typename P::Level L;
      L.qc = g_qc[k]; L.nc = g_nc[k]; L.qr = g_qr[k]; L.nr = g_nr[k]; L.qi = g_qi[k];
      L.qm = g_qm[k]; L.ni = g_ni[k]; L.bm = g_bm[k]; L.qv = g_qv[k]; L.th = g_th[k];
      L.pres = P3_LDG(g_pres + k); L.inv_exner = P3_LDG(g_inv_exner + k);
      L.cf_i = P3_LDG(g_cfi + k); L.cf_l = P3_LDG(g_cfl + k); L.cf_r = P3_LDG(g_cfr + k);
      const Real dz = P3_LDG(g_dz + k), dpres = P3_LDG(g_dpres + k);
      // p3_main_init: p3_main_impl.hpp:21-72
      L.inv_cf_i = 1 / L.cf_i; L.inv_cf_l = 1 / L.cf_l; L.inv_cf_r = 1 / L.cf_r;
      const Real exner = 1 / L.inv_exner;
      const Real T0 = L.th * exner;
      L.qv = P::rmax(L.qv, 0);
      const Real inv_dz = 1 / dz;
      // p3_main_part1: p3_main_impl_part1.hpp:89-182
      L.rho = dpres / dz / C::gravit;
      L.inv_rho = 1 / L.rho;
      err |= P::check_temperature(T0);
      const Real logT0 = std::log(T0);
      const bool cold0 = T0 < Tz;
      Real svp_l = 0;
      bool have_svp_l = false;
      if (!cold0) { svp_l = P::svp_liquid(T0, logT0); have_svp_l = true; }
      L.qv_sat_i = P::qv_from_svp(cold0 ? P::svp_ice(T0, logT0) : svp_l, L.pres);
      L.qv_supersat_i = L.qv / L.qv_sat_i - 1;
      if (cold0 && L.qv_supersat_i >= sp(-0.05)) nuc = true;

      if (L.qc < qsmall) {
        L.qv = L.qv + L.qc;
        L.th = L.th - L.inv_exner * L.qc * latvap * inv_cp;
        L.qc = 0; L.nc = 0;
      } else {
        hyd1 = true;
        if (prescribedCCN)  L.nc = P::rmax(L.nc, a.ro.spa_ccn_to_nc_factor * P3_LDG(g_nccn + k) / L.inv_cf_l);
        else if (predictNc) L.nc = P::rmax(L.nc + P3_LDG(g_nc_nuceat + k) * dt, sp(0.0));
        else                L.nc = C::NCCNST * L.inv_rho;
      }
      if (L.qr < qsmall) {
        L.qv = L.qv + L.qr;
        L.th = L.th - L.inv_exner * L.qr * latvap * inv_cp;
        L.qr = 0; L.nr = 0;
      } else hyd1 = true;
      if (L.qi < qsmall || (L.qi < sp(1.e-8) && L.qv_supersat_i < sp(-0.1))) {
        L.qv = L.qv + L.qi;
        L.th = L.th - L.inv_exner * L.qi * (latvap + latice) * inv_cp;
        L.qi = 0; L.ni = 0; L.qm = 0; L.bm = 0;
      } else hyd1 = true;
      if (L.qi >= qsmall && L.qi < sp(1.e-8) && T0 >= Tz) {
        L.qr = L.qr + L.qi;
        L.th = L.th - L.inv_exner * L.qi * latice * inv_cp;
        L.qi = 0; L.ni = 0; L.qm = 0; L.bm = 0;
      }
      L.T = L.th * exner;
      L.T_init = T0;
      if (L.T < C::T_homogfrz) frz.set(slot);

      // p3_main_part2 (p3_main_impl_part2.hpp:121-127: skip_all)
      Real o_depos = 0, o_ptt = 0, o_nevapr = 0;
      const bool skip_all = (L.qc < qsmall && L.qr < qsmall && L.qi < qsmall && L.T < Tz && L.qv_supersat_i < sp(-0.05));
      if (!skip_all) {
        if (!have_svp_l) svp_l = P::svp_liquid(T0, logT0);
        L.qv_sat_l = P::qv_from_svp(svp_l, L.pres);
        typename P::Part2Out o;
        const bool h = P::part2_level(L, a.tb, a.ro, dt, inv_dt, predictNc, prescribedCCN,
                                      g_qv_prev + k, g_t_prev + k, g_ni_act ? g_ni_act + k : nullptr, o, err);
        if (h) hyd2 = true;
        p2.set(slot);
        o_depos = o.qv2qi_depos_tend; o_ptt = o.precip_total_tend; o_nevapr = o.nevapr;
        s.Vq[k] = o.liq_ice_exchange; s.Vn[k] = o.vap_liq_exchange; s.flux[k] = o.vap_ice_exchange;
      }
      // park the column state
      s.qc[k] = L.qc; s.nc[k] = L.nc; s.qr[k] = L.qr; s.nr[k] = L.nr; s.qi[k] = L.qi;
      s.qm[k] = L.qm; s.ni[k] = L.ni; s.bm[k] = L.bm; s.qv[k] = L.qv; s.th[k] = L.th;
      s.rho[k] = L.rho; s.inv_dz[k] = inv_dz; s.aux[k] = T0;
      // outputs that are final after part2 (zero where part2 did not run: p3_main_impl.hpp:211-217)
      colp(F_QV2QI_DEPOS_TEND)[k] = o_depos;
      colp(F_PRECIP_TOTAL_TEND)[k] = o_ptt;
      colp(F_NEVAPR)[k] = o_nevapr;
      g_plf[k] = 0;
      g_pif[k] = 0;
    }
    if (a.zero_flux_top && lane == (nk % WS)) { g_plf[nk] = 0; g_pif[nk] = 0; }

    nuc = w.any(nuc); hyd1 = w.any(hyd1); hyd2 = w.any(hyd2);
    const bool active = nuc || hyd1;
    int dloc[D_NUM];
    for (int j = 0; j < D_NUM; ++j) dloc[j] = -1;
    dloc[D_QC_NSUB] = dloc[D_QR_NSUB] = dloc[D_QI_NSUB] = 0;
    dloc[D_NUC_POSSIBLE] = nuc; dloc[D_HYDRO_PART1] = hyd1;
    if (active) dloc[D_HYDRO_PART2] = hyd2;

    if (active) {   // history arrays are only touched where part2 ran (p3_main_impl_part2.hpp:468-470)
      Real* lie = colp(F_LIQ_ICE_EXCHANGE); Real* vle = colp(F_VAP_LIQ_EXCHANGE); Real* vie = colp(F_VAP_ICE_EXCHANGE);
      slot = 0;
      for (int k = lane; k < nk; k += WS, ++slot)
        if (p2.get(slot)) { lie[k] = s.Vq[k]; vle[k] = s.Vn[k]; vie[k] = s.flux[k]; }
    }
    w.sync();

    Real precip_liq_surf = 0, precip_ice_surf = 0;
    const bool full = active && hyd2;
    if (full) {
      Real acc;
      sediment<SP_CLOUD>(w, lane, a, col, s, inv_dt, acc, dloc, err);
      if (dloc[D_QC_KTOP] >= 0) precip_liq_surf = acc;               // assignment, p3_cloud_sed_impl.hpp:134
      sediment<SP_RAIN>(w, lane, a, col, s, inv_dt, acc, dloc, err);
      if (dloc[D_QR_KTOP] >= 0) precip_liq_surf += acc;              // p3_rain_sed_impl.hpp:157
      sediment<SP_ICE>(w, lane, a, col, s, inv_dt, acc, dloc, err);
      if (dloc[D_QI_KTOP] >= 0) precip_ice_surf += acc;              // p3_ice_sed_impl.hpp:192
    }

    // ================= pass 3: homogeneous freezing + part3 + write-back ==========================
    Real* const o_rqc = colp(F_DIAG_EFF_RADIUS_QC); Real* const o_rqi = colp(F_DIAG_EFF_RADIUS_QI);
    Real* const o_rqr = colp(F_DIAG_EFF_RADIUS_QR); Real* const o_rho_qi = colp(F_RHO_QI);
    Real* const g_vle = colp(F_VAP_LIQ_EXCHANGE);
    slot = 0;
    for (int k = lane; k < nk; k += WS, ++slot) {
      Real qc = s.qc[k], nc = s.nc[k], qr = s.qr[k], nr = s.nr[k], qi = s.qi[k], qm = s.qm[k],
           ni = s.ni[k], bm = s.bm[k], qv = s.qv[k], th = s.th[k];
      typename P::Part3Out o;
      o.eff_radius_qc = sp(10.e-6); o.eff_radius_qi = sp(25.e-6); o.eff_radius_qr = sp(500.e-6); o.rho_qi = 0;
      if (full) {
        const Real inv_exner = P3_LDG(g_inv_exner + k);
        if (a.ro.do_ice_production)
          P::homogeneous_freezing_level(frz.get(slot), inv_exner, qc, nc, qr, nr, qi, ni, qm, bm, th);
        const Real rho = s.rho[k];
        Real sub_qc, sub_qr;
        P::part3_level(qc, nc, qr, nr, qi, ni, qm, bm, qv, th, rho, 1 / rho, inv_exner,
                       P3_LDG(g_cfl + k), P3_LDG(g_cfr + k), P3_LDG(g_cfi + k), a.tb, a.ro, o, sub_qc, sub_qr);
        if (sub_qc != 0 || sub_qr != 0) {   // vap_liq_exchange read-modify-write, p3_main_impl_part3.hpp:97,127
          Real v = g_vle[k];
          v = v - sub_qc;
          v = v - sub_qr;
          g_vle[k] = v;
        }
      }
      g_qc[k] = qc; g_nc[k] = nc; g_qr[k] = qr; g_nr[k] = nr; g_qi[k] = qi; g_qm[k] = qm;
      g_ni[k] = ni; g_bm[k] = bm; g_qv[k] = qv; g_th[k] = th;
      o_rqc[k] = o.eff_radius_qc; o_rqi[k] = o.eff_radius_qi; o_rqr[k] = o.eff_radius_qr; o_rho_qi[k] = o.rho_qi;
    }
    err = w.or_int(err);
    dloc[D_ERROR] = err;
    if (lane == 0) {
      a.f[F_PRECIP_LIQ_SURF][col] = precip_liq_surf;
      a.f[F_PRECIP_ICE_SURF][col] = precip_ice_surf;
      if (a.diag) {
        int* d = a.diag + col * D_NUM;
        for (int j = 0; j < D_NUM; ++j) d[j] = dloc[j];
      }
    }
    w.sync();   // the shared rows are reused by the next column of this warp
  }
};

}  // namespace p3b200
#endif
