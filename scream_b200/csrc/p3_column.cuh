// p3_column.cuh — the three phases of p3_main for one column on one warp (B200 layout: levels
// strided across the 32 lanes, lane l owns levels l, l+32, ...).
//
// Replaces the body of the "p3 main loop" team lambda (p3/impl/p3_main_impl.hpp:108-325).  The
// reference runs 8 team-parallel phases separated by 15 team_barrier sites inside ONE kernel and
// streams 44 per-column temporaries through a global-memory WorkspaceManager.  Here the column
// passes through three kernels, each shaped for what it does (see p3_kernel.cu):
//
//   phase A  p3_main_init + p3_main_part1 + p3_main_part2 fused per level (all part1->part2
//            temporaries stay in registers): level-parallel, register-heavy, no serial chain;
//   phase B  cloud / rain / ice sedimentation, one launch per species: a serial CFL sub-step loop
//            per column on a REGISTER WINDOW of the levels the hydrometeor can reach — few
//            registers, so many resident warps hide the latency of the dependent chain;
//   phase C  homogeneous_freezing + p3_main_part3, one thread per level.
//
// What travels between the phases (all in arrays the caller already owns, nothing is allocated):
//   - the prognostic state in place (A writes the post-part2 state, B updates the sedimenting
//     windows, C writes the final state);
//   - per-column flags and the sedimentation ranges in the integer diagnostics (D_*);
//   - one per-level scratch value parked in diag_eff_radius_qc, which p3_main only defines at the
//     very end (part3): +-T_atm(init), the sign carrying "T_atm(part1) < T_homogfrz".  Phase B (cloud)
//     needs |.| for acn (p3_main_impl_part1.hpp:108-109), phase C the sign for homogeneous_freezing.
//   Everything else the reference keeps per column (rho, inv_rho, inv_dz, rhofac*, *_incld, dsd
//   parameters) is a pure function of the inputs/prognostics and is re-derived where it is consumed.
//
// Why fusing part1 and part2 is exact: the reference skips part2 for a column iff no level can
// nucleate and no level holds hydrometeors after part1 (p3_main_impl.hpp:237).  For such a column
// every non-skip_all level has zero in-cloud mass and T >= 273.15 K, so all 29 tendencies are exactly
// zero and part2 leaves every prognostic bit-identical; the only visible difference would be the
// three history arrays, whose stores are therefore deferred until the column flags are known.
//
// WP is the warp policy: WarpDev (32 lanes, shuffles/redux) on the GPU, WarpSim (1 lane) in the
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

enum Species { SP_CLOUD = 0, SP_RAIN = 1, SP_ICE = 2 };

constexpr int MAX_NLEV = 256;
constexpr int HIST_ROWS = 3;        // phase A parks the three history outputs of a column in shared memory

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
  int*    diag;            // [ncol][D_NUM], always present (the handle owns one if the caller passes none)
  unsigned int* queue;     // work-queue counters, one per sedimentation launch (species x window class), zeroed before phase A
};

// lane-private bit set indexed by level slot (k / WP::size)
template <int NWORDS>
struct LaneBits {
  unsigned w[NWORDS];
  P3_HD void clear() { for (int i = 0; i < NWORDS; ++i) w[i] = 0u; }
  P3_HD void set(int j) { w[j >> 5] |= (1u << (j & 31)); }
  P3_HD bool get(int j) const { return (w[j >> 5] >> (j & 31)) & 1u; }
};

template <typename Real>
P3_HD Real* col_ptr(const MainArgs<Real>& a, int f, int64_t col) { return a.f[f] + col * a.stride[f]; }

// =================================================================================================
// Phase A: p3_main_init (p3_main_impl.hpp:21-72) + p3_main_part1 (p3_main_impl_part1.hpp:21-185) +
// p3_main_part2 (p3_main_impl_part2.hpp:21-526) for one column; `hist` = HIST_ROWS*nk Reals of scratch.
template <typename Real, typename WP>
struct PhaseA {
  using P = Phys<Real>;
  using C = Cst<Real>;
  static constexpr int WS = WP::size;
  static P3_HD Real sp(double x) { return static_cast<Real>(x); }

  static P3_HD void run(const WP& w, const MainArgs<Real>& a, int64_t col, Real* hist) {
    const int lane = w.lane();
    const int nk = a.nk;
    const Real qsmall = C::QSMALL, Tz = C::Tmelt, latvap = C::LatVap, latice = C::LatIce, inv_cp = C::INV_CP;
    const Real dt = a.dt, inv_dt = 1 / a.dt;
    const bool predictNc = a.predictNc != 0, prescribedCCN = a.prescribedCCN != 0;

    Real* const g_qc = col_ptr(a, F_QC, col); Real* const g_nc = col_ptr(a, F_NC, col);
    Real* const g_qr = col_ptr(a, F_QR, col); Real* const g_nr = col_ptr(a, F_NR, col);
    Real* const g_qi = col_ptr(a, F_QI, col); Real* const g_qm = col_ptr(a, F_QM, col);
    Real* const g_ni = col_ptr(a, F_NI, col); Real* const g_bm = col_ptr(a, F_BM, col);
    Real* const g_qv = col_ptr(a, F_QV, col); Real* const g_th = col_ptr(a, F_TH, col);
    const Real* const g_cfi = col_ptr(a, F_CLD_FRAC_I, col); const Real* const g_cfl = col_ptr(a, F_CLD_FRAC_L, col);
    const Real* const g_cfr = col_ptr(a, F_CLD_FRAC_R, col); const Real* const g_pres = col_ptr(a, F_PRES, col);
    const Real* const g_dz = col_ptr(a, F_DZ, col); const Real* const g_dpres = col_ptr(a, F_DPRES, col);
    const Real* const g_inv_exner = col_ptr(a, F_INV_EXNER, col);
    const Real* const g_qv_prev = col_ptr(a, F_QV_PREV, col); const Real* const g_t_prev = col_ptr(a, F_T_PREV, col);
    const Real* const g_nccn = a.f[F_NCCN] ? col_ptr(a, F_NCCN, col) : nullptr;
    const Real* const g_nc_nuceat = a.f[F_NC_NUCEAT_TEND] ? col_ptr(a, F_NC_NUCEAT_TEND, col) : nullptr;
    const Real* const g_ni_act = a.f[F_NI_ACTIVATED] ? col_ptr(a, F_NI_ACTIVATED, col) : nullptr;
    Real* const g_plf = col_ptr(a, F_PRECIP_LIQ_FLUX, col); Real* const g_pif = col_ptr(a, F_PRECIP_ICE_FLUX, col);
    Real* const g_depos = col_ptr(a, F_QV2QI_DEPOS_TEND, col); Real* const g_ptt = col_ptr(a, F_PRECIP_TOTAL_TEND, col);
    Real* const g_nevapr = col_ptr(a, F_NEVAPR, col);
    Real* const g_stash = col_ptr(a, F_DIAG_EFF_RADIUS_QC, col);
    Real* const h_lie = hist; Real* const h_vle = hist + nk; Real* const h_vie = hist + 2 * nk;

    int err = 0;
    bool nuc = false, hyd1 = false, hyd2 = false;
    // find_top / find_bottom (p3_find_impl.hpp) of the three sedimenting species on the post-part2
    // state, accumulated per lane and reduced once per column
    int lo_c = MAX_NLEV + 1, hi_c = -1, lo_r = MAX_NLEV + 1, hi_r = -1, lo_i = MAX_NLEV + 1, hi_i = -1;
    LaneBits<(MAX_NLEV / WS + 31) / 32> p2;   // per level slot of this lane: the level ran part2
    p2.clear();

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
      const bool frz = L.T < C::T_homogfrz;

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
        h_lie[k] = o.liq_ice_exchange; h_vle[k] = o.vap_liq_exchange; h_vie[k] = o.vap_ice_exchange;
      }
      if (L.qc >= qsmall) { if (k < lo_c) lo_c = k; hi_c = k; }
      if (L.qr >= qsmall) { if (k < lo_r) lo_r = k; hi_r = k; }
      if (L.qi >= qsmall) { if (k < lo_i) lo_i = k; hi_i = k; }
      // the post-part2 state goes straight back to HBM
      g_qc[k] = L.qc; g_nc[k] = L.nc; g_qr[k] = L.qr; g_nr[k] = L.nr; g_qi[k] = L.qi;
      g_qm[k] = L.qm; g_ni[k] = L.ni; g_bm[k] = L.bm; g_qv[k] = L.qv; g_th[k] = L.th;
      // outputs that are final after part2 (zero where part2 did not run: p3_main_impl.hpp:211-217)
      g_depos[k] = o_depos; g_ptt[k] = o_ptt; g_nevapr[k] = o_nevapr;
      g_plf[k] = 0; g_pif[k] = 0;
      g_stash[k] = frz ? -T0 : T0;
    }
    if (a.zero_flux_top && lane == (nk % WS)) { g_plf[nk] = 0; g_pif[nk] = 0; }

    nuc = w.any(nuc); hyd1 = w.any(hyd1); hyd2 = w.any(hyd2);
    const bool active = nuc || hyd1;
    const bool full = active && hyd2;
    // history arrays are only touched where part2 ran (p3_main_impl_part2.hpp:468-470)
    {
      Real* lie = col_ptr(a, F_LIQ_ICE_EXCHANGE, col); Real* vle = col_ptr(a, F_VAP_LIQ_EXCHANGE, col);
      Real* vie = col_ptr(a, F_VAP_ICE_EXCHANGE, col);
      slot = 0;
      for (int k = lane; k < nk; k += WS, ++slot) {
        if (active && p2.get(slot)) { lie[k] = h_lie[k]; vle[k] = h_vle[k]; vie[k] = h_vie[k]; }
      }
    }
    if (!full) {   // the column leaves p3_main here (p3_main_impl.hpp:237,258): part3 outputs keep their defaults
      Real* const o_rqc = col_ptr(a, F_DIAG_EFF_RADIUS_QC, col); Real* const o_rqi = col_ptr(a, F_DIAG_EFF_RADIUS_QI, col);
      Real* const o_rqr = col_ptr(a, F_DIAG_EFF_RADIUS_QR, col); Real* const o_rho_qi = col_ptr(a, F_RHO_QI, col);
      for (int k = lane; k < nk; k += WS) {
        o_rqc[k] = sp(10.e-6); o_rqi[k] = sp(25.e-6); o_rqr[k] = sp(500.e-6); o_rho_qi[k] = 0;
      }
    }
    lo_c = w.min_int(lo_c); hi_c = w.max_int(hi_c);
    lo_r = w.min_int(lo_r); hi_r = w.max_int(hi_r);
    lo_i = w.min_int(lo_i); hi_i = w.max_int(hi_i);
    err = w.or_int(err);
    if (lane == 0) {
      a.f[F_PRECIP_LIQ_SURF][col] = 0;
      a.f[F_PRECIP_ICE_SURF][col] = 0;
      int* d = a.diag + col * D_NUM;
      d[D_NUC_POSSIBLE] = nuc; d[D_HYDRO_PART1] = hyd1; d[D_HYDRO_PART2] = active ? (int)hyd2 : -1;
      const bool c = full && hi_c >= 0, r = full && hi_r >= 0, i = full && hi_i >= 0;
      d[D_QC_KTOP] = c ? lo_c : -1; d[D_QC_KBOT] = c ? hi_c : -1; d[D_QC_NSUB] = 0;
      d[D_QR_KTOP] = r ? lo_r : -1; d[D_QR_KBOT] = r ? hi_r : -1; d[D_QR_NSUB] = 0;
      d[D_QI_KTOP] = i ? lo_i : -1; d[D_QI_KBOT] = i ? hi_i : -1; d[D_QI_NSUB] = 0;
      d[D_ERROR] = err; d[13] = -1; d[14] = -1; d[15] = -1;
    }
    w.sync();   // the history rows are reused by the next column of this warp
  }
};

// =================================================================================================
// Phase B: cloud_sedimentation (p3_cloud_sed_impl.hpp:17-147), rain_sedimentation
// (p3_rain_sed_impl.hpp:39-170) and ice_sedimentation (p3_ice_sed_impl.hpp:69-205) share this CFL
// sub-step loop (generalized_sedimentation + calc_first_order_upwind_step, p3_upwind_impl.hpp:18-129);
// SP selects the fall-speed physics, JW the number of register slots per lane.
//
// Layout: the levels [k_qxtop, nk) that the hydrometeor can reach form a WINDOW anchored at k_qxtop;
// lane l keeps levels k_qxtop + l + j*WS (j = 0..JW-1) entirely in registers for the whole sub-step
// loop (state, rho, 1/rho, 1/dz, cloud fraction and its reciprocal, lazily evaluated density factor).
// A sub-step is then: fall speeds per lane, one warp max (redux) for the Courant number, and per field
// one shuffle for the inflow flux — no shared memory, no block barrier.
//
// Deviations from the reference's operation order, all at rounding level (<= 1-2 ulp per sub-step):
//   - cloud and rain: after the first sub-step the in-cloud values are q * (1/cld_frac) instead of
//     q / cld_frac (ice keeps the division: the rime density qm/bm amplifies last-bit differences);
//   - rain: the mean-size argument of the fall-speed table is (mu_r+1)*cbrt(qr/(c*nr)) instead of
//     (mu_r+1)/cbrt(c*nr/qr), the quotient formed with a 1-ulp reciprocal (Phys::fast_rcp); the lambda
//     limiters are applied on the equivalent bounds with the clamped nr = qr*(lamr^3/c) pre-factored.
// In-cloud values on the FIRST sub-step are what calculate_incloud_mixingratios left at the end of
// part2 (limits included, p3_incloud_mixingratios_impl.hpp:12-69).
template <typename Real, typename WP, int SP, int JW>
struct Sediment {
  using P = Phys<Real>;
  using C = Cst<Real>;
  static constexpr int WS = WP::size;
  static P3_HD Real sp(double x) { return static_cast<Real>(x); }

  // One first-order upwind sub-step of one field held in the register window (kdir = -1: inflow from
  // level k-1).  Slot j of a lane holds level kw + lane + j*WS, so the level above lives in the previous
  // lane of the same slot (one shuffle) or, for lane 0, in the last lane of the previous slot.
  static P3_HD void upwind(const WP& w, int lane, int kw, int k_temp, Real dt_sub, const Real (&V)[JW],
                           Real (&r)[JW], const Real (&rho)[JW], const Real (&inv_rho)[JW],
                           const Real (&inv_dz)[JW], Real (&flux)[JW]) {
    Real prev_last = 0;
#pragma unroll
    for (int j = 0; j < JW; ++j) {
      const int k = kw + lane + j * WS;
      const bool in = k <= k_temp;
      const Real f = in ? V[j] * r[j] * rho[j] : Real(0);
      Real up = w.shfl_up1(f);
      if (lane == 0) up = prev_last;
      if (JW > 1) prev_last = w.shfl_from(f, WS - 1);
      if (in) {
        const Real fluxdiv = (up - f) * inv_dz[j];
        r[j] += fluxdiv * dt_sub * inv_rho[j];
      }
      flux[j] = f;
    }
  }

  static P3_HD void run(const WP& w, const MainArgs<Real>& a, int64_t col, int k_qxtop, int k_qxbot) {
    const int lane = w.lane();
    const int nk = a.nk, kbot = nk - 1;
    const Real qsmall = C::QSMALL;
    constexpr int fq = (SP == SP_CLOUD) ? F_QC : (SP == SP_RAIN) ? F_QR : F_QI;
    constexpr int fn = (SP == SP_CLOUD) ? F_NC : (SP == SP_RAIN) ? F_NR : F_NI;
    constexpr int fcf = (SP == SP_CLOUD) ? F_CLD_FRAC_L : (SP == SP_RAIN) ? F_CLD_FRAC_R : F_CLD_FRAC_I;
    constexpr int dbase = (SP == SP_CLOUD) ? D_QC_KTOP : (SP == SP_RAIN) ? D_QR_KTOP : D_QI_KTOP;
    Real* const g_q = col_ptr(a, fq, col); Real* const g_n = col_ptr(a, fn, col);
    Real* const g_qm = col_ptr(a, F_QM, col); Real* const g_bm = col_ptr(a, F_BM, col);
    const Real* const g_cf = col_ptr(a, fcf, col);
    const Real* const g_cfl = col_ptr(a, F_CLD_FRAC_L, col);
    const Real* const g_dz = col_ptr(a, F_DZ, col); const Real* const g_dpres = col_ptr(a, F_DPRES, col);
    const Real* const g_stash = col_ptr(a, F_DIAG_EFF_RADIUS_QC, col);
    Real* const g_plf = col_ptr(a, F_PRECIP_LIQ_FLUX, col);
    const bool predictNc = a.predictNc != 0;
    const Real inv_dt = 1 / a.dt;

    const int kw = k_qxtop;
    Real q[JW], n[JW], qm[JW], bm[JW], rho[JW], inv_rho[JW], inv_dz[JW], cf[JW], icf[JW], fac[JW], acc[JW];
    // fac: cloud: acn; rain: rhofacr; ice: rhofaci
#pragma unroll
    for (int j = 0; j < JW; ++j) {
      const int k = kw + lane + j * WS;
      if (k < nk) {
        q[j] = g_q[k]; n[j] = g_n[k];
        if (SP == SP_ICE) { qm[j] = g_qm[k]; bm[j] = g_bm[k]; }
        const Real dz = P3_LDG(g_dz + k);
        rho[j] = P3_LDG(g_dpres + k) / dz / C::gravit;   // p3_main_impl_part1.hpp:93
        inv_rho[j] = 1 / rho[j];
        inv_dz[j] = 1 / dz;                               // p3_main_impl.hpp:57
        cf[j] = P3_LDG(g_cf + k);
        icf[j] = 1 / cf[j];
        if (SP == SP_RAIN) acc[j] = g_plf[k + 1];
        if (SP == SP_CLOUD) {   // acn = g*rhow/(18*mu_air(T_atm at init)), p3_main_impl_part1.hpp:108-109
          const Real T0 = std::abs(g_stash[k]);
          const Real dum = sp(1.496e-6) * P::pow1p5(T0) / (T0 + 120);
          fac[j] = C::gravit * C::RHO_H2O / (18 * dum);
        } else if (SP == SP_RAIN) {
          fac[j] = std::pow(C::RHO_1000MB * inv_rho[j], sp(.54));   // rhofacr, p3_main_impl_part1.hpp:105
        } else {
          fac[j] = std::pow(C::RHO_600MB * inv_rho[j], sp(.54));    // rhofaci, p3_main_impl_part1.hpp:106
        }
      } else {
        q[j] = 0; n[j] = 0; rho[j] = 1; inv_rho[j] = 1; inv_dz[j] = 0; cf[j] = 1; icf[j] = 1; fac[j] = 0;
        if (SP == SP_ICE) { qm[j] = 0; bm[j] = 0; }
        if (SP == SP_RAIN) acc[j] = 0;
      }
    }

    int nsub = 0, err = 0;
    Real dt_left = a.dt, prt_local = 0;
    bool first = true;
    // rain: bounds on dum1 = (mu_r+1)/lamr equivalent to the lambda limiters of get_rain_dsd2
    const Real mu_r = a.ro.constant_mu_rain;
    const Real mass_to_d3 = C::CONS1 * (mu_r + 3) * (mu_r + 2) * (mu_r + 1);
    const Real lamr_max = (mu_r + sp(1.)) * sp(1.e+5), lamr_min = (mu_r + sp(1.)) * 500;
    const Real c_min = 1 / lamr_max, c_max = 1 / lamr_min;    // bounds on 1/lamr
    // nr = lamr^3 qr / mass_to_d3 at the two limiters (p3_dsd2_impl.hpp:116-119)
    const Real nr_per_q_min = lamr_min * lamr_min * lamr_min / mass_to_d3, nr_per_q_max = lamr_max * lamr_max * lamr_max / mass_to_d3;

    while (dt_left > C::dt_left_tol) {
      const int k_temp = (k_qxbot == kbot) ? k_qxbot : k_qxbot + 1;
      Real Vq[JW], Vn[JW];
      Real Co_loc = 0;
#pragma unroll
      for (int j = 0; j < JW; ++j) {
        const int k = kw + lane + j * WS;
        Real vq = 0, vn = 0;
        if (k <= k_qxbot) {
          if (SP == SP_CLOUD) {
            Real qc_i, nc_i;
            if (first) P::incloud_liq(q[j], n[j], icf[j], qc_i, nc_i);
            else { qc_i = q[j] * icf[j]; nc_i = n[j] * icf[j]; }
            if (qc_i > qsmall) {
              Real mu_c, lamc, d0, d1;
              P::cloud_dsd2(qc_i, nc_i, rho[j], mu_c, lamc, d0, d1, false);
              n[j] = nc_i * cf[j];
              const Real acn = fac[j];
              // bcn = 2: pow(lamc, bcn) = lamc*lamc exactly for IEEE pow with an integral exponent 2
              const Real dum = 1 / (lamc * lamc);
              // Gamma(4+bcn+mu)/Gamma(mu+4) = (mu+5)(mu+4), Gamma(1+bcn+mu)/Gamma(mu+1) = (mu+2)(mu+1):
              // kept as tgamma quotients to follow the reference's rounding
              vq = acn * std::tgamma(4 + C::bcn + mu_c) * dum / std::tgamma(mu_c + 4);
              if (predictNc) vn = acn * std::tgamma(1 + C::bcn + mu_c) * dum / std::tgamma(mu_c + 1);
            }
          } else if (SP == SP_RAIN) {
            Real qr_i, nr_i;
            if (first) P::incloud_rain(q[j], n[j], icf[j], qr_i, nr_i);
            else { qr_i = q[j] * icf[j]; nr_i = n[j] * icf[j]; }
            if (qr_i > qsmall) {
              // get_rain_dsd2 + compute_rain_fall_velocity (p3_dsd2_impl.hpp:76-127, p3_rain_sed_impl.hpp:17-34)
              const Real nr_lim = P::rmax(nr_i, C::NSMALL);
              Real c = std::cbrt(qr_i * P::fast_rcp(mass_to_d3 * nr_lim));      // 1/lamr
              nr_i = nr_lim;
              if (c > c_max) { nr_i = qr_i * nr_per_q_min; c = c_max; }         // lamr < lammin
              else if (c < c_min) { nr_i = qr_i * nr_per_q_max; c = c_min; }    // lamr > lammax
              typename P::Loc3 t3;
              P::lookup3_from_dum1(mu_r, (mu_r + 1) * c, t3);
              vq = P::apply_table3(a.tb.vm, t3) * fac[j];
              vn = P::apply_table3(a.tb.vn, t3) * fac[j];
              n[j] = nr_i * cf[j];
            }
          } else {
            Real qi_i, ni_i, qm_i, bm_i;
            if (first) P::incloud_ice(q[j], n[j], qm[j], bm[j], icf[j], 1 / P3_LDG(g_cfl + k), qi_i, ni_i, qm_i, bm_i);
            else { qi_i = q[j] / cf[j]; ni_i = n[j] / cf[j]; qm_i = qm[j] / cf[j]; bm_i = bm[j] / cf[j]; }   // exact: rho_rime = qm/bm is ill-conditioned
            if (qi_i > qsmall) {
              ni_i = P::rmax(ni_i, C::NSMALL);
              const Real rhop = P::calc_bulk_rho_rime(qi_i, qm_i, bm_i, a.ro);
              qm[j] = qm_i * cf[j];
              bm[j] = bm_i * cf[j];
              typename P::IceLoc ti;
              P::lookup_ice(qi_i, ni_i, qm_i, rhop, ti);
              Real tv_ni_fallspd, tv_qi_fallspd, tv_ni_lammax, tv_ni_lammin;
              P::apply_table_ice2(0, a.tb.ice, ti, tv_ni_fallspd, tv_qi_fallspd);
              P::apply_table_ice2(6, a.tb.ice, ti, tv_ni_lammax, tv_ni_lammin);
              ni_i = P::rmin(ni_i, tv_ni_lammax * ni_i);
              ni_i = P::rmax(ni_i, tv_ni_lammin * ni_i);
              n[j] = ni_i * cf[j];
              vq = a.ro.ice_sedimentation_factor * tv_qi_fallspd * fac[j];
              vn = a.ro.ice_sedimentation_factor * tv_ni_fallspd * fac[j];
            }
          }
          const Real Co = vq * dt_left * inv_dz[j];
          if (Co > Co_loc) Co_loc = Co;
        }
        Vq[j] = vq; Vn[j] = vn;
      }
      const Real Co_max = w.max_nonneg(Co_loc);
      // generalized_sedimentation: p3_upwind_impl.hpp:97-129
      if (!(Co_max >= 0)) err |= 4;
      const int tmpint1 = static_cast<int>(Co_max + 1);
      const Real dt_sub = dt_left / tmpint1;
      Real flux[JW], ftmp[JW];
      if (SP == SP_ICE) {
        upwind(w, lane, kw, k_temp, dt_sub, Vq, bm, rho, inv_rho, inv_dz, ftmp);
        upwind(w, lane, kw, k_temp, dt_sub, Vq, qm, rho, inv_rho, inv_dz, ftmp);
      }
      if (SP != SP_CLOUD || predictNc) upwind(w, lane, kw, k_temp, dt_sub, Vn, n, rho, inv_rho, inv_dz, ftmp);
      upwind(w, lane, kw, k_temp, dt_sub, Vq, q, rho, inv_rho, inv_dz, flux);
#pragma unroll
      for (int j = 0; j < JW; ++j) {
        const int k = kw + lane + j * WS;
        if (k == kbot && k_qxbot == kbot) prt_local += flux[j] * dt_sub;
        // precip_liq_flux(k+1) += flux_qx(k) for k in [k_qxtop, new k_qxbot], p3_rain_sed_impl.hpp:139-153
        if (SP == SP_RAIN && k <= k_temp) acc[j] = acc[j] + flux[j];
      }
      if (k_qxbot != kbot) k_qxbot += 1;
      dt_left -= dt_sub;
      first = false;
      ++nsub;
    }
    // write back the part of the window the front reached
#pragma unroll
    for (int j = 0; j < JW; ++j) {
      const int k = kw + lane + j * WS;
      if (k <= k_qxbot) {
        g_q[k] = q[j]; g_n[k] = n[j];
        if (SP == SP_ICE) { g_qm[k] = qm[j]; g_bm[k] = bm[j]; }
        if (SP == SP_RAIN) g_plf[k + 1] = acc[j];
      }
    }
    const Real prt_accum = w.shfl_from(prt_local, (kbot - kw) % WS);   // only the lane owning kbot accumulated
    const Real surf = prt_accum * C::INV_RHO_H2O * inv_dt;
    err = w.or_int(err);
    if (lane == 0) {
      // launches run in stream order cloud -> rain -> ice, so these read-modify-writes never race
      if (SP == SP_CLOUD) a.f[F_PRECIP_LIQ_SURF][col] = surf;                 // assignment, p3_cloud_sed_impl.hpp:134
      if (SP == SP_RAIN)  a.f[F_PRECIP_LIQ_SURF][col] += surf;                // p3_rain_sed_impl.hpp:157
      if (SP == SP_ICE)   a.f[F_PRECIP_ICE_SURF][col] += surf;                // p3_ice_sed_impl.hpp:192
      int* d = a.diag + col * D_NUM;
      d[dbase + 1] = k_qxbot; d[dbase + 2] = nsub;
      if (err) d[D_ERROR] |= err;
    }
  }
};

// =================================================================================================
// Phase C: homogeneous_freezing (p3_ice_sed_impl.hpp:210-270) + p3_main_part3
// (p3_main_impl_part3.hpp:21-204) for ONE level of a column that went through sedimentation
// (columns that left p3_main earlier got their part3 defaults in phase A).  Prognostics are only
// stored where part3 changed them: on most levels it is the identity.
template <typename Real>
struct PhaseC {
  using P = Phys<Real>;
  using C = Cst<Real>;
  static P3_HD Real sp(double x) { return static_cast<Real>(x); }
  template <class T> static P3_HD void put(T* p, T v_new, T v_old) { if (!(v_new == v_old)) *p = v_new; }

  static P3_HD void level(const MainArgs<Real>& a, int64_t col, int k) {
    Real* const g_qc = col_ptr(a, F_QC, col) + k; Real* const g_nc = col_ptr(a, F_NC, col) + k;
    Real* const g_qr = col_ptr(a, F_QR, col) + k; Real* const g_nr = col_ptr(a, F_NR, col) + k;
    Real* const g_qi = col_ptr(a, F_QI, col) + k; Real* const g_qm = col_ptr(a, F_QM, col) + k;
    Real* const g_ni = col_ptr(a, F_NI, col) + k; Real* const g_bm = col_ptr(a, F_BM, col) + k;
    Real* const g_stash = col_ptr(a, F_DIAG_EFF_RADIUS_QC, col) + k;
    const Real qc0 = *g_qc, nc0 = *g_nc, qr0 = *g_qr, nr0 = *g_nr, qi0 = *g_qi, qm0 = *g_qm, ni0 = *g_ni, bm0 = *g_bm;
    typename P::Part3Out o;
    o.eff_radius_qc = sp(10.e-6); o.eff_radius_qi = sp(25.e-6); o.eff_radius_qr = sp(500.e-6); o.rho_qi = 0;
    if (qc0 == 0 && qr0 == 0 && qi0 == 0) {
      // nothing to freeze, nothing to clip into vapour (qv + 0, th - x*0 are identities): part3 only
      // zeroes the number/rime fields of the absent species (p3_main_impl_part3.hpp:98-100,128-130,190-196,201)
      put(g_nc, Real(0), nc0); put(g_nr, Real(0), nr0); put(g_ni, Real(0), ni0);
      put(g_qm, Real(0), qm0); put(g_bm, Real(0), bm0);
    } else {
      Real* const g_qv = col_ptr(a, F_QV, col) + k; Real* const g_th = col_ptr(a, F_TH, col) + k;
      const Real qv0 = *g_qv, th0 = *g_th;
      Real qc = qc0, nc = nc0, qr = qr0, nr = nr0, qi = qi0, qm = qm0, ni = ni0, bm = bm0, qv = qv0, th = th0;
      const bool frz = *g_stash < 0;
      const Real inv_exner = P3_LDG(col_ptr(a, F_INV_EXNER, col) + k);
      if (a.ro.do_ice_production)
        P::homogeneous_freezing_level(frz, inv_exner, qc, nc, qr, nr, qi, ni, qm, bm, th);
      // rho and the cloud fractions are only consumed where the species exists
      Real rho = 1, cf_l = 1, cf_r = 1, cf_i = 1;
      if (qc >= C::QSMALL || qi >= C::QSMALL)
        rho = P3_LDG(col_ptr(a, F_DPRES, col) + k) / P3_LDG(col_ptr(a, F_DZ, col) + k) / C::gravit;
      if (qc >= C::QSMALL) cf_l = P3_LDG(col_ptr(a, F_CLD_FRAC_L, col) + k);
      if (qr >= C::QSMALL) cf_r = P3_LDG(col_ptr(a, F_CLD_FRAC_R, col) + k);
      if (qi >= C::QSMALL) cf_i = P3_LDG(col_ptr(a, F_CLD_FRAC_I, col) + k);
      Real sub_qc, sub_qr;
      P::part3_level(qc, nc, qr, nr, qi, ni, qm, bm, qv, th, rho, 1 / rho, inv_exner, cf_l, cf_r, cf_i, a.tb, a.ro, o, sub_qc, sub_qr);
      if (sub_qc != 0 || sub_qr != 0) {   // vap_liq_exchange read-modify-write, p3_main_impl_part3.hpp:97,127
        Real* const g_vle = col_ptr(a, F_VAP_LIQ_EXCHANGE, col) + k;
        Real v = *g_vle;
        v = v - sub_qc;
        v = v - sub_qr;
        *g_vle = v;
      }
      put(g_qc, qc, qc0); put(g_nc, nc, nc0); put(g_qr, qr, qr0); put(g_nr, nr, nr0); put(g_qi, qi, qi0);
      put(g_qm, qm, qm0); put(g_ni, ni, ni0); put(g_bm, bm, bm0); put(g_qv, qv, qv0); put(g_th, th, th0);
    }
    *g_stash = o.eff_radius_qc;
    col_ptr(a, F_DIAG_EFF_RADIUS_QI, col)[k] = o.eff_radius_qi;
    col_ptr(a, F_DIAG_EFF_RADIUS_QR, col)[k] = o.eff_radius_qr;
    col_ptr(a, F_RHO_QI, col)[k] = o.rho_qi;
  }
};

}  // namespace p3b200
#endif
