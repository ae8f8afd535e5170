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

// Kernel arguments (by value in param space).
template <typename Real>
struct MainArgs {
  Real*   f[F_NUM];
  int64_t stride[F_NUM];
  int64_t ustride;         // common column stride of every 2-D field except the two flux arrays (0: strides differ)
  int64_t ustride_flux;    // common column stride of precip_liq_flux / precip_ice_flux
  int     ncol, nk;
  Real    dt;
  int     predictNc, prescribedCCN, zero_flux_top;
  int     zero_history;    // P3B200Infra::zero_history: this call zero-fills the three history arrays (the classification pass writes the zeros)
  Tables<Real>  tb;
  Runtime<Real> ro;
  int*    diag;            // [ncol][D_NUM], always present (the handle owns one if the caller passes none)
  unsigned int* queue;     // work-queue counters: one per sedimentation launch (species x window class) and per part2 group, zeroed before phase A
  // Optional p3_preamble fused into the classification pass (p3b200_main_preamble): the field-manager inputs of
  // eamxx_p3_process_interface.hpp:62-129.  pre_pmid == nullptr: no preamble.  The arrays the preamble writes are the
  // p3_main fields themselves (mixing ratios in place, inv_exner, th, the three cloud fractions, dz, qv_prev).
  const Real *pre_pmid, *pre_pd, *pre_pdd, *pre_T, *pre_cft;
  // phase A scratch owned by the handle (p3_abi.cpp), valid between the classification pass and the column finish:
  unsigned short* wl_idx;  // [ntiles][A_TILE_LEVELS] per tile: tile-local level indices (col_in_tile*nk + k) of the levels part2 has to run on, ordered by process group, index order inside a group
  int*            wl_off;  // [ntiles][8]: start of each process group's segment in the tile's list; [6] = number of entries
  unsigned char*  lvlflags;// [ncol*nk] per level after part2: LF_* bits (species present, hydrometeors, error bits); 0 on skip_all levels
};

template <typename Real>
P3_HD Real* col_ptr(const MainArgs<Real>& a, int f, int64_t col) { return a.f[f] + col * a.stride[f]; }

// Column addressing of the per-level kernels.  UNI = all 2-D fields share one column stride (what EAMxx and the
// test bridge hand over unless a field is a sub-view of a bundled tracer array): one 64-bit multiply per level
// instead of one per field (address arithmetic was ~13% of the instructions of the first kernels).
template <bool UNI, typename Real>
struct ColAddr {
  int64_t col, off, off_flux;
  P3_HD ColAddr(const MainArgs<Real>& a, int64_t c)
      : col(c), off(UNI ? c * a.ustride : 0), off_flux(UNI ? c * a.ustride_flux : 0) {}
  template <int F> P3_HD Real* p(const MainArgs<Real>& a) const {
    if (UNI) return a.f[F] + ((F == F_PRECIP_LIQ_FLUX || F == F_PRECIP_ICE_FLUX) ? off_flux : off);
    return a.f[F] + col * a.stride[F];
  }
};

// =================================================================================================
// Phase A: p3_main_init (p3_main_impl.hpp:21-72) + p3_main_part1 (p3_main_impl_part1.hpp:21-185) +
// p3_main_part2 (p3_main_impl_part2.hpp:21-526).  All three are level-local, so phase A is written per
// LEVEL and the kernel is free to choose which lane runs which level:
//
//   pass 1 (every level, one thread per level, coalesced along the column): init + the part of part1
//          that decides the column flags (nucleationPossible, hydrometeorsPresent) and the level's
//          skip_all gate (p3_main_impl_part2.hpp:121-127).  A skip_all level is COMPLETE after part1 —
//          its post-part1 state is stored here from 7 loads.  Any other level is only classified: the
//          pass returns a sort key (which species are present, temperature regime) and stores nothing.
//   pass 2 (the classified levels, ordered by process group = temperature regime x ice present, in index order
//          inside a group): init + part1 + part2 from the level's untouched inputs, stores, and the
//          per-column reductions (hydrometeorsPresent after part2, find_top/find_bottom of the three
//          sedimenting species) into a small per-column accumulator.
//
// The two passes are separate kernels (p3_kernel.cu): the classification pass is a streaming kernel at the
// register budget of a streaming kernel; part2 runs as one launch per PROCESS GROUP (temperature regime x ice
// present, Phys::part2_level<GRP>) on the group-ordered work lists the classification pass leaves in global memory,
// each group compiled to the code and registers its regime needs.  What a level reports back to its column
// (species present after part2 for find_top/find_bottom, hydrometeorsPresent, error bits) is one byte per level
// (LF_*); a third, tiny kernel folds the bytes of a column into the integer diagnostics.
enum ColFlag { CF_NUC = 1, CF_HYD1 = 2, CF_ERR_SHIFT = 8 };
enum LevelFlag { LF_QC = 1, LF_QR = 2, LF_QI = 4, LF_HYD = 8, LF_ERR_SHIFT = 4 };
constexpr int A_NKEYS = 32;          // sort keys: (tc << 3) | (ice << 2) | (rain << 1) | cloud, tc in {0,1,2}: 24 used
constexpr int A_NGROUPS = 6;         // process group = key >> 2 = 2*tc + ice (Phys::part2_level<GRP>)
constexpr int A_OFF_STRIDE = 8;      // ints per tile in MainArgs::wl_off

template <typename Real>
struct PhaseA {
  using P = Phys<Real>;
  using C = Cst<Real>;
  static P3_HD Real sp(double x) { return static_cast<Real>(x); }
  // bit pattern of +0.0 (a -0.0 is stored: only the value the memory already holds may be skipped)
  static P3_HD bool is_pos_zero(Real x) { return x == 0 && !std::signbit(x); }
  static P3_HD bool same_bits(Real a, Real b) { return (a == b && std::signbit(a) == std::signbit(b)) || (a != a && b != b); }

  // ---- pass 1 -----------------------------------------------------------------------------------
  // m_*: what memory holds for the five values this pass may have to store (with the fused preamble these are the wet
  // mixing ratios and an undefined th: every value the preamble changes is stored)
  struct In1 { Real qc, qr, qi, qv, th, inv_exner, pres, m_qc, m_qr, m_qi, m_qv; bool th_in_memory; };
  struct Pre1 { Real pmid, pd, pdd, T, cft, cft_above, qv_prev; };     // PRE: the preamble's own loads of one level
  template <bool UNI, bool PRE = false>
  static P3_HD void pass1_load(const MainArgs<Real>& a, int64_t col, int k, In1& in, Pre1* pre = nullptr) {
    const ColAddr<UNI, Real> cp(a, col);
    in.qc = cp.template p<F_QC>(a)[k]; in.qr = cp.template p<F_QR>(a)[k]; in.qi = cp.template p<F_QI>(a)[k];
    in.qv = cp.template p<F_QV>(a)[k];
    in.pres = P3_LDG(cp.template p<F_PRES>(a) + k);
    in.m_qc = in.qc; in.m_qr = in.qr; in.m_qi = in.qi; in.m_qv = in.qv;
    if (PRE) {
      const int64_t o = col * a.ustride + k;        // the fused preamble requires the common column stride
      pre->pmid = P3_LDG(a.pre_pmid + o); pre->pd = P3_LDG(a.pre_pd + o); pre->pdd = P3_LDG(a.pre_pdd + o);
      pre->T = P3_LDG(a.pre_T + o); pre->cft = P3_LDG(a.pre_cft + o);
      pre->cft_above = (k != 0) ? P3_LDG(a.pre_cft + o - 1) : Real(0);
      pre->qv_prev = cp.template p<F_QV_PREV>(a)[k];
      in.th = 0; in.inv_exner = 1; in.th_in_memory = false;
    } else {
      in.th = cp.template p<F_TH>(a)[k];
      in.inv_exner = P3_LDG(cp.template p<F_INV_EXNER>(a) + k);
      in.th_in_memory = true;
    }
  }

  // p3_preamble of one level (eamxx_p3_process_interface.hpp:62-129; helper formulas
  // share/util/scream_common_physics_functions_impl.hpp:88-97, 115-118, 214-222, 375-379, 400-409), statement for
  // statement what p3_preamble_kernel (p3_prepost.cu) does: dz from the wet qv, wet -> dry of the four mixing ratios
  // this pass reads and of qv_prev, exner / theta, the cloud fractions with the max-overlap rain fraction.  Stores the
  // six arrays only the preamble defines; the dry qc/qr/qi/qv and th travel on in `in`.
  template <bool UNI>
  static P3_HD void preamble_level(const MainArgs<Real>& a, int64_t col, int k, const Pre1& pr, In1& in) {
    const ColAddr<UNI, Real> cp(a, col);
    const Real c1 = -Real(1) + Real(1) / C::ep_2;
    const Real T_virtual = pr.T * (Real(1) + c1 * in.qv);
    cp.template p<F_DZ>(a)[k] = (C::RD / C::gravit) * pr.pd * T_virtual / pr.pmid;
    // div_z: a clear level's mixing ratios are exactly 0, and a zero numerator sends the fp64 division out of line
    in.qv = P::div_z(in.qv * pr.pd, pr.pdd); in.qc = P::div_z(in.qc * pr.pd, pr.pdd);
    in.qr = P::div_z(in.qr * pr.pd, pr.pdd); in.qi = P::div_z(in.qi * pr.pd, pr.pdd);
    cp.template p<F_QV_PREV>(a)[k] = P::div_z(pr.qv_prev * pr.pd, pr.pdd);
    const Real ex = P::x_pow(pr.pmid / C::P0, C::RD * C::INV_CP);
    in.inv_exner = Real(1.0) / ex;
    in.th = pr.T / ex;
    cp.template p<F_INV_EXNER>(a)[k] = in.inv_exner;
    const Real mincld = Real(0.0001);
    const Real cf = pr.cft < mincld ? mincld : pr.cft;
    cp.template p<F_CLD_FRAC_L>(a)[k] = cf; cp.template p<F_CLD_FRAC_I>(a)[k] = cf;
    cp.template p<F_CLD_FRAC_R>(a)[k] = (k != 0 && pr.cft_above > cf) ? pr.cft_above : cf;
  }
  // ... and of the five mixing ratios the classification pass does not read, on a level part2 will run on
  template <bool UNI>
  static P3_HD void preamble_rest_of_level(const MainArgs<Real>& a, int64_t col, int k, const Pre1& pr) {
    const ColAddr<UNI, Real> cp(a, col);
    Real* const p[5] = {cp.template p<F_NC>(a) + k, cp.template p<F_NR>(a) + k, cp.template p<F_QM>(a) + k,
                        cp.template p<F_NI>(a) + k, cp.template p<F_BM>(a) + k};
    Real v[5];
#pragma unroll
    for (int i = 0; i < 5; ++i) v[i] = *p[i];
#pragma unroll
    for (int i = 0; i < 5; ++i) *p[i] = P::div_z(v[i] * pr.pd, pr.pdd);
  }

  // Single-precision estimate of qv / qv_sat_i(T0, pres) on a cold level (relative error < 1e-4: fp32
  // rounding of the Murphy-Koop ice exponent, |terms| <= 40, plus the fast exp/log).  Pass 1 only needs the
  // three DECISIONS the reference takes on qv_supersat_i (>= -0.05: nucleationPossible and skip_all; < -0.1:
  // small-ice clipping); when the estimate is further than 2e-3 from every threshold that matters the decision
  // is already certain and the fp64 evaluation is skipped.  Returns false when the exact value is needed.
  static P3_HD bool supersat_decided(Real T0, Real pres, Real qv0, bool small_ice, Real& supersat_i) {
    if constexpr (sizeof(Real) == 8) {
      const float Tf = (float)T0;
#if defined(__CUDA_ARCH__)
      const float e = __expf(9.550426f - __fdividef(5723.265f, Tf) + 3.53068f * __logf(Tf) - 0.00728332f * Tf);
      // approximate division (2 ulp, inside the estimate's error budget): the IEEE one took its out-of-line slow path
      // on every level with qv = 0
      const float ratio = __fdividef((float)qv0 * fmaxf((float)pres, 1.e-3f), (float)C::ep_2 * e);
#else
      const float e = std::exp(9.550426f - 5723.265f / Tf + 3.53068f * std::log(Tf) - 0.00728332f * Tf);
      const float ratio = (float)qv0 * fmaxf((float)pres, 1.e-3f) / ((float)C::ep_2 * e);
#endif
      const bool clear95 = fabsf(ratio - 0.95f) > 2.e-3f;
      const bool clear90 = !small_ice || fabsf(ratio - 0.90f) > 2.e-3f;
      if (clear95 && clear90 && ratio < 1.e30f) {     // NaN/inf estimates fall through to the exact path
        supersat_i = (Real)ratio - 1;
        return true;
      }
      return false;
    } else {
      return false;
    }
  }

  // Returns -1 when the level is finished (skip_all; state stored), else its sort key in [0, A_NKEYS).
  // `flags` receives CF_NUC / CF_HYD1 / error bits of this level (to be OR-ed into the column's accumulator).
  template <bool UNI, bool PRE = false>
  static P3_HD int pass1_level(const MainArgs<Real>& a, int64_t col, int k, const In1& in, int& flags, const Pre1* pre = nullptr) {
    const ColAddr<UNI, Real> cp(a, col);
    const Real qsmall = C::QSMALL, Tz = C::Tmelt, latvap = C::LatVap, latice = C::LatIce, inv_cp = C::INV_CP;
    if (a.zero_history) {   // the caller's zero-fill of eamxx_p3_run.cpp:27-29, folded into this pass
      cp.template p<F_LIQ_ICE_EXCHANGE>(a)[k] = 0; cp.template p<F_VAP_LIQ_EXCHANGE>(a)[k] = 0; cp.template p<F_VAP_ICE_EXCHANGE>(a)[k] = 0;
    }
    Real qc = in.qc, qr = in.qr, qi = in.qi, qv = in.qv, th = in.th;
    const Real inv_exner = in.inv_exner;
    // p3_main_init
    const Real exner = 1 / inv_exner;
    const Real T0 = th * exner;
    qv = P::rmax(qv, 0);
    int err = P::check_temperature(T0);
    const bool cold0 = T0 < Tz;
    // qv_supersat_i (p3_main_impl_part1.hpp:95-101) is consumed by nucleationPossible on cold levels, by the
    // small-ice clipping when qi is in [qsmall, 1e-8) and by skip_all; evaluated only where one of them asks
    const Real qv0 = qv;
    const bool small_ice = qi >= qsmall && qi < sp(1.e-8);
    Real supersat_i = 0;
    bool have_sat = false;
    auto eval_sat = [&]() {
      const Real logT0 = P::x_log(T0);
      const Real svp = cold0 ? P::svp_ice(T0, logT0) : P::svp_liquid(T0, logT0);
      supersat_i = qv0 / P::qv_from_svp(svp, in.pres) - 1;
      have_sat = true;
    };
    if (cold0) {
      if (supersat_decided(T0, in.pres, qv0, small_ice, supersat_i)) have_sat = true;
      else eval_sat();
    } else if (small_ice) {
      eval_sat();
    }
    flags = 0;
    if (cold0 && supersat_i >= sp(-0.05)) flags |= CF_NUC;
    if (qc < qsmall) { qv = qv + qc; th = th - inv_exner * qc * latvap * inv_cp; qc = 0; } else flags |= CF_HYD1;
    if (qr < qsmall) { qv = qv + qr; th = th - inv_exner * qr * latvap * inv_cp; qr = 0; } else flags |= CF_HYD1;
    if (qi < qsmall || (qi < sp(1.e-8) && supersat_i < sp(-0.1))) {
      qv = qv + qi; th = th - inv_exner * qi * (latvap + latice) * inv_cp; qi = 0;
    } else flags |= CF_HYD1;
    if (qi >= qsmall && qi < sp(1.e-8) && T0 >= Tz) {
      qr = qr + qi; th = th - inv_exner * qi * latice * inv_cp; qi = 0;
    }
    const Real T = th * exner;
    const bool has_c = qc >= qsmall, has_r = qr >= qsmall, has_i = qi >= qsmall;
    bool skip_all = !has_c && !has_r && !has_i && T < Tz;
    if (skip_all) {
      if (!have_sat) eval_sat();          // T0 >= Tmelt but the clipped level is colder: rare
      skip_all = supersat_i < sp(-0.05);
    }
    flags |= err << CF_ERR_SHIFT;
    if (!skip_all) {
      // What part2 leaves at its default on most levels is written HERE, with the stores of the neighbouring
      // skip_all levels (full lines): the three tendencies and the two flux arrays at zero, and the parked
      // +-T_atm(init).  The part2 launches then store a tendency only where it is non-zero: a lone store into a
      // line that is not in L2 costs a full-line fill from DRAM.
      cp.template p<F_QV2QI_DEPOS_TEND>(a)[k] = 0; cp.template p<F_PRECIP_TOTAL_TEND>(a)[k] = 0; cp.template p<F_NEVAPR>(a)[k] = 0;
      cp.template p<F_PRECIP_LIQ_FLUX>(a)[k] = 0; cp.template p<F_PRECIP_ICE_FLUX>(a)[k] = 0;
      cp.template p<F_DIAG_EFF_RADIUS_QC>(a)[k] = (T < C::T_homogfrz) ? -T0 : T0;
      if (PRE) {   // part2 reads its inputs from memory: the preamble's results of this level have to be there
        cp.template p<F_QC>(a)[k] = in.qc; cp.template p<F_QR>(a)[k] = in.qr; cp.template p<F_QI>(a)[k] = in.qi;
        cp.template p<F_QV>(a)[k] = in.qv; cp.template p<F_TH>(a)[k] = in.th;
        preamble_rest_of_level<UNI>(a, col, k, *pre);
      }
      // temperature regime of Phys::part2_level<GRP>: 0 T > Tmelt, 1 T_rainfrz < T <= Tmelt, 2 T <= T_rainfrz
      const int tc = (T <= C::T_rainfrz) ? 2 : (T <= Tz) ? 1 : 0;
      return (tc << 3) | (int(has_i) << 2) | (int(has_r) << 1) | int(has_c);
    }
    // the level is complete: every hydrometeor was clipped into vapour (p3_main_impl_part1.hpp:123-175),
    // part2 does not run (p3_main_impl.hpp:211-217: its outputs stay zero)
    // the five values this pass has loaded are stored only where their bits change (a clear level stays clear)
    if (!is_pos_zero(in.m_qc)) cp.template p<F_QC>(a)[k] = 0;
    if (!is_pos_zero(in.m_qr)) cp.template p<F_QR>(a)[k] = 0;
    if (!is_pos_zero(in.m_qi)) cp.template p<F_QI>(a)[k] = 0;
    cp.template p<F_NC>(a)[k] = 0; cp.template p<F_NR>(a)[k] = 0;
    cp.template p<F_QM>(a)[k] = 0; cp.template p<F_NI>(a)[k] = 0; cp.template p<F_BM>(a)[k] = 0;
    if (!same_bits(qv, in.m_qv)) cp.template p<F_QV>(a)[k] = qv;
    if (!in.th_in_memory || !same_bits(th, in.th)) cp.template p<F_TH>(a)[k] = th;
    cp.template p<F_QV2QI_DEPOS_TEND>(a)[k] = 0; cp.template p<F_PRECIP_TOTAL_TEND>(a)[k] = 0; cp.template p<F_NEVAPR>(a)[k] = 0;
    cp.template p<F_PRECIP_LIQ_FLUX>(a)[k] = 0; cp.template p<F_PRECIP_ICE_FLUX>(a)[k] = 0;
    cp.template p<F_DIAG_EFF_RADIUS_QC>(a)[k] = (T < C::T_homogfrz) ? -T0 : T0;
    return -1;
  }

  // ---- pass 2 -----------------------------------------------------------------------------------
  // `col_active` = nucleationPossible || hydrometeorsPresent of the column (known after pass 1).  GRP = the level's
  // process group (key >> 2) or -1.  Returns the level's LF_* byte.
  template <bool UNI, int GRP>
  static P3_HD int pass2_level(const MainArgs<Real>& a, int64_t col, int k, bool col_active) {
    const ColAddr<UNI, Real> cp(a, col);
    const Real qsmall = C::QSMALL, Tz = C::Tmelt, latvap = C::LatVap, latice = C::LatIce, inv_cp = C::INV_CP;
    const Real dt = a.dt, inv_dt = 1 / a.dt;
    const bool predictNc = a.predictNc != 0, prescribedCCN = a.prescribedCCN != 0;
    Real* const g_qc = cp.template p<F_QC>(a) + k; Real* const g_nc = cp.template p<F_NC>(a) + k;
    Real* const g_qr = cp.template p<F_QR>(a) + k; Real* const g_nr = cp.template p<F_NR>(a) + k;
    Real* const g_qi = cp.template p<F_QI>(a) + k; Real* const g_qm = cp.template p<F_QM>(a) + k;
    Real* const g_ni = cp.template p<F_NI>(a) + k; Real* const g_bm = cp.template p<F_BM>(a) + k;
    Real* const g_qv = cp.template p<F_QV>(a) + k; Real* const g_th = cp.template p<F_TH>(a) + k;

    typename P::Level L;
    L.qc = *g_qc; L.nc = *g_nc; L.qr = *g_qr; L.nr = *g_nr; L.qi = *g_qi;
    L.qm = *g_qm; L.ni = *g_ni; L.bm = *g_bm; L.qv = *g_qv; L.th = *g_th;
    // which of the eight hydrometeor values are +0 in memory: an absent species stays absent on most levels, and a store
    // of the +0 that is already there only dirties a sector the part2 launches then write back (they are bound by
    // exactly that traffic)
    const unsigned zero_in = (is_pos_zero(L.qc) ? 1u : 0u) | (is_pos_zero(L.nc) ? 2u : 0u) | (is_pos_zero(L.qr) ? 4u : 0u) |
                             (is_pos_zero(L.nr) ? 8u : 0u) | (is_pos_zero(L.qi) ? 16u : 0u) | (is_pos_zero(L.qm) ? 32u : 0u) |
                             (is_pos_zero(L.ni) ? 64u : 0u) | (is_pos_zero(L.bm) ? 128u : 0u);
    L.pres = P3_LDG(cp.template p<F_PRES>(a) + k); L.inv_exner = P3_LDG(cp.template p<F_INV_EXNER>(a) + k);
    L.cf_i = P3_LDG(cp.template p<F_CLD_FRAC_I>(a) + k); L.cf_l = P3_LDG(cp.template p<F_CLD_FRAC_L>(a) + k);
    L.cf_r = P3_LDG(cp.template p<F_CLD_FRAC_R>(a) + k);
    const Real dz = P3_LDG(cp.template p<F_DZ>(a) + k), dpres = P3_LDG(cp.template p<F_DPRES>(a) + k);
    // p3_main_init: p3_main_impl.hpp:21-72
    L.inv_cf_i = 1 / L.cf_i; L.inv_cf_l = 1 / L.cf_l; L.inv_cf_r = 1 / L.cf_r;
    const Real exner = 1 / L.inv_exner;
    const Real T0 = L.th * exner;
    L.qv = P::rmax(L.qv, 0);
    // p3_main_part1: p3_main_impl_part1.hpp:89-182
    L.rho = dpres / dz / C::gravit;
    L.inv_rho = 1 / L.rho;
    int err = P::check_temperature(T0);
    const Real logT0 = P::x_log(T0);
    const bool cold0 = T0 < Tz;
    const Real svp_l = P::svp_liquid(T0, logT0);
    L.qv_sat_l = P::qv_from_svp(svp_l, L.pres);
    L.qv_sat_i = cold0 ? P::qv_from_svp(P::svp_ice(T0, logT0), L.pres) : L.qv_sat_l;
    L.qv_supersat_i = L.qv / L.qv_sat_i - 1;

    if (L.qc < qsmall) {
      L.qv = L.qv + L.qc;
      L.th = L.th - L.inv_exner * L.qc * latvap * inv_cp;
      L.qc = 0; L.nc = 0;
    } else {
      if (prescribedCCN)  L.nc = P::rmax(L.nc, a.ro.spa_ccn_to_nc_factor * P3_LDG(cp.template p<F_NCCN>(a) + k) / L.inv_cf_l);
      else if (predictNc) L.nc = P::rmax(L.nc + P3_LDG(cp.template p<F_NC_NUCEAT_TEND>(a) + k) * dt, sp(0.0));
      else                L.nc = C::NCCNST * L.inv_rho;
    }
    if (L.qr < qsmall) {
      L.qv = L.qv + L.qr;
      L.th = L.th - L.inv_exner * L.qr * latvap * inv_cp;
      L.qr = 0; L.nr = 0;
    }
    if (L.qi < qsmall || (L.qi < sp(1.e-8) && L.qv_supersat_i < sp(-0.1))) {
      L.qv = L.qv + L.qi;
      L.th = L.th - L.inv_exner * L.qi * (latvap + latice) * inv_cp;
      L.qi = 0; L.ni = 0; L.qm = 0; L.bm = 0;
    }
    if (L.qi >= qsmall && L.qi < sp(1.e-8) && T0 >= Tz) {
      L.qr = L.qr + L.qi;
      L.th = L.th - L.inv_exner * L.qi * latice * inv_cp;
      L.qi = 0; L.ni = 0; L.qm = 0; L.bm = 0;
    }
    L.T = L.th * exner;
    L.T_init = T0;
    const bool frz = L.T < C::T_homogfrz;

    // p3_main_part2 (this level is not skip_all: pass 1 decided that on the same values)
    typename P::Part2Out o;
    const Real* const g_ni_act = a.f[F_NI_ACTIVATED] ? cp.template p<F_NI_ACTIVATED>(a) + k : nullptr;
    const bool h = P::template part2_level<GRP>(L, a.tb, a.ro, dt, inv_dt, predictNc, prescribedCCN,
                                                cp.template p<F_QV_PREV>(a) + k, cp.template p<F_T_PREV>(a) + k, g_ni_act, o, err);
    // what find_top / find_bottom (p3_find_impl.hpp) and hydrometeorsPresent need from this level
    const int lf = (L.qc >= qsmall ? LF_QC : 0) | (L.qr >= qsmall ? LF_QR : 0) | (L.qi >= qsmall ? LF_QI : 0) |
                   (h ? LF_HYD : 0) | (err << LF_ERR_SHIFT);
    auto put = [&](Real* p, Real v, unsigned bit) { if (!((zero_in & bit) && is_pos_zero(v))) *p = v; };
    put(g_qc, L.qc, 1u); put(g_nc, L.nc, 2u); put(g_qr, L.qr, 4u); put(g_nr, L.nr, 8u); put(g_qi, L.qi, 16u);
    put(g_qm, L.qm, 32u); put(g_ni, L.ni, 64u); put(g_bm, L.bm, 128u);
    *g_qv = L.qv; *g_th = L.th;
    // the classification pass wrote zero tendencies, zero fluxes and the parked +-T_atm(init) for this level:
    // only a tendency whose bit pattern is not +0 needs a store
    if (!is_pos_zero(o.qv2qi_depos_tend))  cp.template p<F_QV2QI_DEPOS_TEND>(a)[k] = o.qv2qi_depos_tend;
    if (!is_pos_zero(o.precip_total_tend)) cp.template p<F_PRECIP_TOTAL_TEND>(a)[k] = o.precip_total_tend;
    if (!is_pos_zero(o.nevapr))            cp.template p<F_NEVAPR>(a)[k] = o.nevapr;
    (void)frz;
    // history arrays are only touched where part2 ran (p3_main_impl_part2.hpp:468-470), and part2 only runs
    // for columns that can nucleate or hold hydrometeors after part1 (p3_main_impl.hpp:237); for any other
    // column all tendencies of this level are exactly zero and the prognostics above are unchanged.  With
    // zero_history the classification pass has zero-filled the three arrays: a +0 needs no store either.
    if (col_active) {
      if (!a.zero_history || !is_pos_zero(o.liq_ice_exchange)) cp.template p<F_LIQ_ICE_EXCHANGE>(a)[k] = o.liq_ice_exchange;
      if (!a.zero_history || !is_pos_zero(o.vap_liq_exchange)) cp.template p<F_VAP_LIQ_EXCHANGE>(a)[k] = o.vap_liq_exchange;
      if (!a.zero_history || !is_pos_zero(o.vap_ice_exchange)) cp.template p<F_VAP_ICE_EXCHANGE>(a)[k] = o.vap_ice_exchange;
    }
    return lf;
  }

  // run-time dispatch on the process group (host build; the kernels are instantiated per group)
  template <bool UNI>
  static P3_HD int pass2_level_any(const MainArgs<Real>& a, int64_t col, int k, bool col_active, int grp) {
    switch (grp) {
      case 0: return pass2_level<UNI, 0>(a, col, k, col_active);
      case 1: return pass2_level<UNI, 1>(a, col, k, col_active);
      case 2: return pass2_level<UNI, 2>(a, col, k, col_active);
      case 3: return pass2_level<UNI, 3>(a, col, k, col_active);
      case 4: return pass2_level<UNI, 4>(a, col, k, col_active);
      case 5: return pass2_level<UNI, 5>(a, col, k, col_active);
      default: return pass2_level<UNI, -1>(a, col, k, col_active);
    }
  }

  // ---- per column, after the classification pass: what part1 decided (fl = OR of the levels' pass-1 flags)
  static P3_HD void column_begin(const MainArgs<Real>& a, int64_t col, int fl) {
    a.f[F_PRECIP_LIQ_SURF][col] = 0;
    a.f[F_PRECIP_ICE_SURF][col] = 0;
    if (a.zero_flux_top) { col_ptr(a, F_PRECIP_LIQ_FLUX, col)[a.nk] = 0; col_ptr(a, F_PRECIP_ICE_FLUX, col)[a.nk] = 0; }
    int* d = a.diag + col * D_NUM;
    d[D_NUC_POSSIBLE] = (fl & CF_NUC) != 0; d[D_HYDRO_PART1] = (fl & CF_HYD1) != 0;
    d[D_ERROR] = fl >> CF_ERR_SHIFT;
  }

  // ---- per column, after part2: hydrometeorsPresent and the sedimentation ranges (find_top / find_bottom,
  // p3_find_impl.hpp) from the levels' LF_* bytes into the integer diagnostics
  static P3_HD void column_finish(const MainArgs<Real>& a, int64_t col, const unsigned char* lf) {
    int* d = a.diag + col * D_NUM;
    int any = 0, lo_c = MAX_NLEV + 1, hi_c = -1, lo_r = MAX_NLEV + 1, hi_r = -1, lo_i = MAX_NLEV + 1, hi_i = -1;
    for (int k = 0; k < a.nk; ++k) {
      const int f = lf[k];
      any |= f;
      if (f & LF_QC) { if (k < lo_c) lo_c = k; hi_c = k; }
      if (f & LF_QR) { if (k < lo_r) lo_r = k; hi_r = k; }
      if (f & LF_QI) { if (k < lo_i) lo_i = k; hi_i = k; }
    }
    const bool active = d[D_NUC_POSSIBLE] || d[D_HYDRO_PART1];
    const bool hyd2 = (any & LF_HYD) != 0;
    const bool full = active && hyd2;
    d[D_HYDRO_PART2] = active ? (int)hyd2 : -1;
    const bool c = full && hi_c >= 0, r = full && hi_r >= 0, i = full && hi_i >= 0;
    d[D_QC_KTOP] = c ? lo_c : -1; d[D_QC_KBOT] = c ? hi_c : -1; d[D_QC_NSUB] = 0;
    d[D_QR_KTOP] = r ? lo_r : -1; d[D_QR_KBOT] = r ? hi_r : -1; d[D_QR_NSUB] = 0;
    d[D_QI_KTOP] = i ? lo_i : -1; d[D_QI_KBOT] = i ? hi_i : -1; d[D_QI_NSUB] = 0;
    d[D_ERROR] |= any >> LF_ERR_SHIFT; d[13] = -1; d[14] = -1; d[15] = -1;
  }

  // Host build only (tests/hostsim): the passes for one column in the order the kernels run them.
  static inline void run_column_host(const MainArgs<Real>& a, int64_t col) {
    int keys[MAX_NLEV];
    unsigned char lf[MAX_NLEV];
    int flags = 0;
    for (int k = 0; k < a.nk; ++k) {
      int fl = 0;
      In1 in;
      pass1_load<false>(a, col, k, in);
      keys[k] = pass1_level<false>(a, col, k, in, fl);
      flags |= fl;
      lf[k] = 0;
    }
    column_begin(a, col, flags);
    const bool active = (flags & (CF_NUC | CF_HYD1)) != 0;
    for (int key = 0; key < A_NKEYS; ++key)
      for (int k = 0; k < a.nk; ++k)
        if (keys[k] == key) lf[k] = (unsigned char)pass2_level_any<false>(a, col, k, active, key >> 2);
    column_finish(a, col, lf);
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
      if (JW > 1 && kw + j * WS > k_temp) { flux[j] = 0; prev_last = 0; continue; }   // warp-uniform: slot below the front
      const bool in = k <= k_temp;
      const Real f = in ? V[j] * r[j] * rho[j] : Real(0);
      Real up = w.shfl_up1(f);
      if (lane == 0) up = prev_last;
      if (JW > 1) prev_last = w.shfl_from(f, WS - 1);
      // select instead of a branch (lanes outside the range compute a value nobody keeps)
      const Real fluxdiv = (up - f) * inv_dz[j];
      const Real r_new = r[j] + fluxdiv * dt_sub * inv_rho[j];
      r[j] = in ? r_new : r[j];
      flux[j] = f;
    }
  }

  // rain_tab (rain only): the 300 x (vm, vn) slice of Tables::vmn for this run's mu_r column, or nullptr
  // RFAST (rain only): the caller guarantees rain_tab != nullptr, and the generic table path is compiled out of the loop
  template <bool RFAST = false>
  static P3_HD void run(const WP& w, const MainArgs<Real>& a, int64_t col, int k_qxtop, int k_qxbot,
                        const Real* rain_tab = nullptr) {
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
    // all loads of the window first (up to 7 per slot in flight), then the divisions: the setup of a tall window
    // was one exposed DRAM latency per slot and field
    Real dzw[JW], dpw[JW], t0w[JW], cflw[JW];   // cflw: ice only, the liquid cloud fraction of bm_incld (first sub-step)
#pragma unroll
    for (int j = 0; j < JW; ++j) {
      const int k = kw + lane + j * WS;
      const bool in = k < nk;
      q[j] = in ? g_q[k] : Real(0); n[j] = in ? g_n[k] : Real(0);
      if (SP == SP_ICE) { qm[j] = in ? g_qm[k] : Real(0); bm[j] = in ? g_bm[k] : Real(0); }
      dzw[j] = in ? P3_LDG(g_dz + k) : Real(1);
      dpw[j] = in ? P3_LDG(g_dpres + k) : Real(1);
      cf[j] = in ? P3_LDG(g_cf + k) : Real(1);
      if (SP == SP_RAIN) acc[j] = in ? g_plf[k + 1] : Real(0);
      if (SP == SP_CLOUD) t0w[j] = in ? std::abs(g_stash[k]) : Real(1);
      if (SP == SP_ICE) cflw[j] = in ? P3_LDG(g_cfl + k) : Real(1);
    }
#pragma unroll
    for (int j = 0; j < JW; ++j) {
      const int k = kw + lane + j * WS;
      if (k < nk) {
        const Real dz = dzw[j];
        rho[j] = dpw[j] / dz / C::gravit;                 // p3_main_impl_part1.hpp:93
        inv_rho[j] = 1 / rho[j];
        inv_dz[j] = 1 / dz;                               // p3_main_impl.hpp:57
        icf[j] = 1 / cf[j];
        if (SP == SP_CLOUD) {   // acn = g*rhow/(18*mu_air(T_atm at init)), p3_main_impl_part1.hpp:108-109
          const Real T0 = t0w[j];
          const Real dum = sp(1.496e-6) * P::pow1p5(T0) / (T0 + 120);
          fac[j] = C::gravit * C::RHO_H2O / (18 * dum);
        } else if (SP == SP_RAIN) {
          fac[j] = P::x_pow(C::RHO_1000MB * inv_rho[j], sp(.54));   // rhofacr, p3_main_impl_part1.hpp:105
        } else {
          fac[j] = -1;    // rhofaci (p3_main_impl_part1.hpp:106) is evaluated when the level first holds ice
        }
      } else {
        rho[j] = 1; inv_rho[j] = 1; inv_dz[j] = 0; icf[j] = 1; fac[j] = 0;
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
    // the mu_r axis of the fall-speed tables (p3_table3_impl.hpp:52-62) is the same on every level: when its
    // interpolation weight is exactly 0 (integral mu_r) the bilinear form reduces bit for bit to the linear
    // interpolation along the size axis, read from the (vm, vn)-interleaved slice `rain_tab`
    // (the caller passes rain_tab only when that holds: p3_sediment checks the weight once per CTA)
#if defined(P3B200_EXACT_MATH)
    const bool rain_fast = false;       // exact-order build: get_rain_dsd2 + lookup + apply_table as the reference calls them
    (void)rain_tab;
#else
    const bool rain_fast = (SP == SP_RAIN) && (RFAST || rain_tab != nullptr);
#endif

    // rain, fast-table path: the in-cloud values of the NEXT sub-step are carried in registers (what
    // calculate_incloud_mixingratios left after part2 for the first one, q*(1/cld_frac) afterwards), and the
    // fall-speed block is straight-line code on selects — in the first version 31 of the 256 instructions of a
    // sub-step were BSSY/BRA/BSYNC and 33 were register moves of the branch joins
    Real q_in[JW], n_in[JW];
    if (SP == SP_RAIN) {
#pragma unroll
      for (int j = 0; j < JW; ++j) P::incloud_rain(q[j], n[j], icf[j], q_in[j], n_in[j]);
    }

    while (dt_left > C::dt_left_tol) {
      const int k_temp = (k_qxbot == kbot) ? k_qxbot : k_qxbot + 1;
      Real Vq[JW], Vn[JW];
      Real Co_loc = 0;
#pragma unroll
      for (int j = 0; j < JW; ++j) {
        const int k = kw + lane + j * WS;
        Real vq = 0, vn = 0;
        if (JW > 1 && kw + j * WS > k_qxbot) { Vq[j] = 0; Vn[j] = 0; continue; }   // warp-uniform: the front has not reached this slot
        if (SP == SP_RAIN && rain_fast) {
          const bool act = (k <= k_qxbot) && (q_in[j] > qsmall);
          const Real qs = act ? q_in[j] : Real(1), ns = act ? n_in[j] : Real(1);
          // get_rain_dsd2 + compute_rain_fall_velocity (p3_dsd2_impl.hpp:76-127, p3_rain_sed_impl.hpp:17-34)
          const Real nr_lim = P::rmax(ns, C::NSMALL);
          Real c = P::cbrt_quot(qs, mass_to_d3 * nr_lim);                  // 1/lamr
          const bool lam_lo = c > c_max, lam_hi = c < c_min;               // lamr < lammin / lamr > lammax
          const Real nr_i = lam_lo ? qs * nr_per_q_min : (lam_hi ? qs * nr_per_q_max : nr_lim);
          c = lam_lo ? c_max : (lam_hi ? c_min : c);
          // lookup (p3_table3_impl.hpp:17-65) with selects instead of the two-regime branch
          const Real dum1 = (mu_r + 1) * c;
          const bool sm = dum1 <= sp(195.e-6);
          const Real x = dum1 * sp(1.e6);
          Real r = sm ? (x + 5) * sp(0.1) : (x - 195) * (C::THIRD * sp(0.1)) + 20;
          r = P::rmax(r, sm ? sp(1.0) : sp(20.0));
          r = P::rmin(r, sm ? sp(20.0) : sp(300.0));
          int d = static_cast<int>(r);
          d = P::imin(P::imax(d, sm ? 1 : 20), sm ? 20 : 299);
          const Real wgt = r - Real(d);
          const typename P::Real2 t0 = P::ld2_any(rain_tab + 2 * (d - 1)), t1 = P::ld2_any(rain_tab + 2 * d);
          vq = act ? (t0.x + wgt * (t1.x - t0.x)) * fac[j] : Real(0);
          vn = act ? (t0.y + wgt * (t1.y - t0.y)) * fac[j] : Real(0);
          n[j] = act ? nr_i * cf[j] : n[j];
          const Real Co = vq * dt_left * inv_dz[j];
          Co_loc = (Co > Co_loc) ? Co : Co_loc;
          Vq[j] = vq; Vn[j] = vn;
          continue;
        }
        if (SP == SP_ICE) {
          // rhofaci for the whole slot the first time any of its levels holds ice (one pow per slot and column, not
          // one per sub-step as the front advances level by level); all lanes of the warp take part in the vote
          if (w.any(fac[j] < 0 && k <= k_qxbot && q[j] > 0)) { if (fac[j] < 0) fac[j] = P::x_pow(C::RHO_600MB * inv_rho[j], sp(.54)); }
        }
        if (k <= k_qxbot) {
          if (SP == SP_CLOUD) {
            Real qc_i, nc_i;
            if (first) P::incloud_liq(q[j], n[j], icf[j], qc_i, nc_i);
#if defined(P3B200_EXACT_MATH)
            else { qc_i = q[j] / cf[j]; nc_i = n[j] / cf[j]; }       // p3_cloud_sed_impl.hpp:124-125
#else
            else { qc_i = q[j] * icf[j]; nc_i = n[j] * icf[j]; }
#endif
            if (qc_i > qsmall) {
              Real mu_c, lamc, d0, d1;
              P::template cloud_dsd2<true>(qc_i, nc_i, rho[j], mu_c, lamc, d0, d1, false);
              n[j] = nc_i * cf[j];
              const Real acn = fac[j];
              // bcn = 2: pow(lamc, bcn) = lamc*lamc exactly for IEEE pow with an integral exponent 2
              const Real dum = 1 / P::pow_bcn(lamc);
              // bcn = 2: Gamma(4+bcn+mu)/Gamma(mu+4) = (mu+5)(mu+4) and Gamma(1+bcn+mu)/Gamma(mu+1) = (mu+2)(mu+1)
              // exactly; the products are within 1 ulp of the true quotient, closer than the quotient of two
              // library tgamma values the reference forms (4 tgamma calls were 40 % of this kernel's instructions)
#if defined(P3B200_EXACT_MATH)
              vq = acn * P::x_tgamma(4 + C::bcn + mu_c) * dum / P::x_tgamma(mu_c + 4);
              if (predictNc) vn = acn * P::x_tgamma(1 + C::bcn + mu_c) * dum / P::x_tgamma(mu_c + 1);
#else
              vq = acn * ((mu_c + 5) * (mu_c + 4)) * dum;
              if (predictNc) vn = acn * ((mu_c + 2) * (mu_c + 1)) * dum;
#endif
            }
          } else if (SP == SP_RAIN) {
            Real qr_i = q_in[j], nr_i = n_in[j];
#if defined(P3B200_EXACT_MATH)
            if (qr_i > qsmall) {
              // compute_rain_fall_velocity (p3_rain_sed_impl.hpp:17-34) statement by statement
              Real mu_r_k = 0, lamr;
              P::rain_dsd2(qr_i, nr_i, mu_r_k, lamr, a.ro);
              typename P::Loc3 t3;
              P::lookup3(mu_r_k, lamr, t3);
              vq = P::apply_table3(a.tb.vm, t3) * fac[j];
              vn = P::apply_table3(a.tb.vn, t3) * fac[j];
              n[j] = nr_i * cf[j];
            }
#else
            if (qr_i > qsmall) {
              // get_rain_dsd2 + compute_rain_fall_velocity (p3_dsd2_impl.hpp:76-127, p3_rain_sed_impl.hpp:17-34)
              const Real nr_lim = P::rmax(nr_i, C::NSMALL);
              Real c = P::cbrt_quot(qr_i, mass_to_d3 * nr_lim);                // 1/lamr
              nr_i = nr_lim;
              if (c > c_max) { nr_i = qr_i * nr_per_q_min; c = c_max; }         // lamr < lammin
              else if (c < c_min) { nr_i = qr_i * nr_per_q_max; c = c_min; }    // lamr > lammax
              if (rain_fast) {
                // lookup (p3_table3_impl.hpp:17-65) with selects instead of the two-regime branch
                const Real dum1 = (mu_r + 1) * c;
                const bool sm = dum1 <= sp(195.e-6);
                const Real x = dum1 * sp(1.e6);
                Real r = sm ? (x + 5) * sp(0.1) : (x - 195) * (C::THIRD * sp(0.1)) + 20;
                r = P::rmax(r, sm ? sp(1.0) : sp(20.0));
                r = P::rmin(r, sm ? sp(20.0) : sp(300.0));
                int d = static_cast<int>(r);
                d = P::imin(P::imax(d, sm ? 1 : 20), sm ? 20 : 299);
                const Real wgt = r - Real(d);
                const typename P::Real2 t0 = P::ld2_any(rain_tab + 2 * (d - 1)), t1 = P::ld2_any(rain_tab + 2 * d);
                vq = (t0.x + wgt * (t1.x - t0.x)) * fac[j];
                vn = (t0.y + wgt * (t1.y - t0.y)) * fac[j];
              } else {
                typename P::Loc3 t3;
                P::lookup3_from_dum1(mu_r, (mu_r + 1) * c, t3);
                vq = P::apply_table3(a.tb.vm, t3) * fac[j];
                vn = P::apply_table3(a.tb.vn, t3) * fac[j];
              }
              n[j] = nr_i * cf[j];
            }
#endif
          } else {
            Real qi_i, ni_i, qm_i, bm_i;
            if (first) P::incloud_ice(q[j], n[j], qm[j], bm[j], icf[j], 1 / cflw[j], qi_i, ni_i, qm_i, bm_i);
            else {   // exact quotients: rho_rime = qm/bm is ill-conditioned.  div_z: most levels of a window are empty
              qi_i = P::div_z(q[j], cf[j]); ni_i = P::div_z(n[j], cf[j]); qm_i = P::div_z(qm[j], cf[j]); bm_i = P::div_z(bm[j], cf[j]);
            }
            if (qi_i > qsmall) {
              ni_i = P::rmax(ni_i, C::NSMALL);
              const Real rhop = P::calc_bulk_rho_rime(qi_i, qm_i, bm_i, a.ro);
              qm[j] = qm_i * cf[j];
              bm[j] = bm_i * cf[j];
              typename P::IceLoc ti;
              P::template lookup_ice<true>(qi_i, ni_i, qm_i, rhop, ti);
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
      const Real dt_sub = P::div_by_count(dt_left, tmpint1);
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
        const Real prt_new = prt_local + flux[j] * dt_sub;
        prt_local = (k == kbot && k_qxbot == kbot) ? prt_new : prt_local;
        // precip_liq_flux(k+1) += flux_qx(k) for k in [k_qxtop, new k_qxbot], p3_rain_sed_impl.hpp:139-153
        // (flux is exactly 0 outside that range and acc >= 0, so the add needs no guard)
#if defined(P3B200_EXACT_MATH)
        if (SP == SP_RAIN) { acc[j] = acc[j] + flux[j]; q_in[j] = q[j] / cf[j]; n_in[j] = n[j] / cf[j]; }   // p3_rain_sed_impl.hpp:133-134
#else
        if (SP == SP_RAIN) { acc[j] = acc[j] + flux[j]; q_in[j] = q[j] * icf[j]; n_in[j] = n[j] * icf[j]; }
#endif
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
      // The three species' launches may run concurrently (p3_kernel.cu: launch_main).  precip_liq_surf is zeroed by
      // phase A, assigned by cloud (p3_cloud_sed_impl.hpp:134) and incremented by rain (p3_rain_sed_impl.hpp:157):
      // (0 + c) + r == (0 + r) + c bit for bit (IEEE addition commutes, 0 + x is exact), so both ADD atomically.
      if (SP == SP_CLOUD) WP::atomic_add(a.f[F_PRECIP_LIQ_SURF] + col, surf);
      if (SP == SP_RAIN)  WP::atomic_add(a.f[F_PRECIP_LIQ_SURF] + col, surf);
      if (SP == SP_ICE)   a.f[F_PRECIP_ICE_SURF][col] += surf;                // p3_ice_sed_impl.hpp:192 (only ice writes it)
      int* d = a.diag + col * D_NUM;
      d[dbase + 1] = k_qxbot; d[dbase + 2] = nsub;
      if (err) WP::atomic_or(d + D_ERROR, err);
    }
  }

  // ---- ice on tall windows: the same sub-step loop with the window in a per-warp array instead of registers ----
  // run() unrolls the JW slots of a lane: for ice at JW = 4 that is an 83 KB loop body, and with every warp of an SM at
  // a different place in it the kernel waits on instruction fetch (stall "no instruction" 5 warps per issue, issue
  // slots 36 % busy; profiles/r02h_sediment_random_ic_ncu_summary.txt).  Here the slots are visited by a ROLLED loop:
  // one copy of the fall-speed block and one of the upwind block, the state of slot j read from and written back to
  // `st` (shared memory on the device: RS_NUM arrays of JW*WS values, element j*WS + lane: conflict-free).  Level by
  // level the arithmetic is the one of run(): same expressions, same order; the four upwind sweeps of a sub-step are
  // independent of each other and become one sweep over the slots that carries four inflow values.
  enum RolledState { RS_Q, RS_N, RS_QM, RS_BM, RS_RHO, RS_IRHO, RS_IDZ, RS_CF, RS_FAC, RS_VQ, RS_VN, RS_NUM };
  static constexpr int rolled_state_size = RS_NUM * JW * WS;

  static P3_HD void run_rolled(const WP& w, const MainArgs<Real>& a, int64_t col, int k_qxtop, int k_qxbot, Real* st) {
    static_assert(SP == SP_ICE, "the rolled window is built for ice_sedimentation");
    constexpr int NS = JW * WS;
    const int lane = w.lane();
    const int nk = a.nk, kbot = nk - 1;
    const Real qsmall = C::QSMALL;
    Real* const g_q = col_ptr(a, F_QI, col); Real* const g_n = col_ptr(a, F_NI, col);
    Real* const g_qm = col_ptr(a, F_QM, col); Real* const g_bm = col_ptr(a, F_BM, col);
    const Real* const g_cf = col_ptr(a, F_CLD_FRAC_I, col);
    const Real* const g_cfl = col_ptr(a, F_CLD_FRAC_L, col);
    const Real* const g_dz = col_ptr(a, F_DZ, col); const Real* const g_dpres = col_ptr(a, F_DPRES, col);
    const Real inv_dt = 1 / a.dt;
    const int kw = k_qxtop;
    const int nslot = (nk - kw + WS - 1) / WS;          // slots that hold a level (<= JW by the launch's window class)
    Real* const s_q = st + RS_Q * NS + lane;   Real* const s_n = st + RS_N * NS + lane;
    Real* const s_qm = st + RS_QM * NS + lane; Real* const s_bm = st + RS_BM * NS + lane;
    Real* const s_rho = st + RS_RHO * NS + lane; Real* const s_irho = st + RS_IRHO * NS + lane;
    Real* const s_idz = st + RS_IDZ * NS + lane; Real* const s_cf = st + RS_CF * NS + lane;
    Real* const s_fac = st + RS_FAC * NS + lane; Real* const s_vq = st + RS_VQ * NS + lane;
    Real* const s_vn = st + RS_VN * NS + lane;
    // A slot is loaded (7 loads, 3 divisions per level) when the front first reaches it, not before: a window spans
    // everything down to the surface, but the front advances one level per sub-step — a column with ice aloft and a
    // single sub-step used to set up three slots to use one.  A lane only ever touches its own elements of `st`: no
    // synchronisation anywhere.
    int nready = 0;
    auto load_slot = [&](int j) {
      const int k = kw + lane + j * WS, o = j * WS;
      const bool in = k < nk;
      const Real q = in ? g_q[k] : Real(0), n = in ? g_n[k] : Real(0);
      const Real qm = in ? g_qm[k] : Real(0), bm = in ? g_bm[k] : Real(0);
      const Real dz = in ? P3_LDG(g_dz + k) : Real(1), dp = in ? P3_LDG(g_dpres + k) : Real(1);
      const Real cf = in ? P3_LDG(g_cf + k) : Real(1);
      s_q[o] = q; s_n[o] = n; s_qm[o] = qm; s_bm[o] = bm; s_cf[o] = cf;
      if (in) {
        const Real rho = dp / dz / C::gravit;             // p3_main_impl_part1.hpp:93
        s_rho[o] = rho; s_irho[o] = 1 / rho; s_idz[o] = 1 / dz;   // p3_main_impl.hpp:57
        s_fac[o] = -1;    // rhofaci (p3_main_impl_part1.hpp:106) is evaluated when the level first holds ice
      } else {
        s_rho[o] = 1; s_irho[o] = 1; s_idz[o] = 0; s_fac[o] = 0;
      }
    };

    int nsub = 0, err = 0;
    Real dt_left = a.dt, prt_local = 0;
    bool first = true;
    while (dt_left > C::dt_left_tol) {
      const int k_temp = (k_qxbot == kbot) ? k_qxbot : k_qxbot + 1;
#pragma unroll 1
      while (nready < nslot && kw + nready * WS <= k_temp) { load_slot(nready); ++nready; }   // warp-uniform
      Real Co_loc = 0;
#pragma unroll 1
      for (int j = 0; j < nslot && kw + j * WS <= k_qxbot; ++j) {     // warp-uniform: slots the front has reached
        const int k = kw + lane + j * WS, o = j * WS;
        Real vq = 0, vn = 0;
        Real fac = s_fac[o];
        const Real q = s_q[o], inv_rho = s_irho[o];
        // rhofaci for the whole slot the first time any of its levels holds ice; all lanes take part in the vote
        if (w.any(fac < 0 && k <= k_qxbot && q > 0)) {
          if (fac < 0) { fac = P::x_pow(C::RHO_600MB * inv_rho, sp(.54)); s_fac[o] = fac; }
        }
        if (k <= k_qxbot) {
          const Real cf = s_cf[o];
          const Real n = s_n[o], qm = s_qm[o], bm = s_bm[o];
          Real qi_i, ni_i, qm_i, bm_i;
          if (first) P::incloud_ice(q, n, qm, bm, 1 / cf, 1 / P3_LDG(g_cfl + k), qi_i, ni_i, qm_i, bm_i);
          else {   // exact quotients: rho_rime = qm/bm is ill-conditioned.  div_z: most levels of a window are empty
            qi_i = P::div_z(q, cf); ni_i = P::div_z(n, cf); qm_i = P::div_z(qm, cf); bm_i = P::div_z(bm, cf);
          }
          if (qi_i > qsmall) {
            ni_i = P::rmax(ni_i, C::NSMALL);
            const Real rhop = P::calc_bulk_rho_rime(qi_i, qm_i, bm_i, a.ro);
            s_qm[o] = qm_i * cf;
            s_bm[o] = bm_i * cf;
            typename P::IceLoc ti;
            P::template lookup_ice<true>(qi_i, ni_i, qm_i, rhop, ti);
            Real tv_ni_fallspd, tv_qi_fallspd, tv_ni_lammax, tv_ni_lammin;
            P::apply_table_ice2(0, a.tb.ice, ti, tv_ni_fallspd, tv_qi_fallspd);
            P::apply_table_ice2(6, a.tb.ice, ti, tv_ni_lammax, tv_ni_lammin);
            ni_i = P::rmin(ni_i, tv_ni_lammax * ni_i);
            ni_i = P::rmax(ni_i, tv_ni_lammin * ni_i);
            s_n[o] = ni_i * cf;
            vq = a.ro.ice_sedimentation_factor * tv_qi_fallspd * fac;
            vn = a.ro.ice_sedimentation_factor * tv_ni_fallspd * fac;
          }
          const Real Co = vq * dt_left * s_idz[o];
          if (Co > Co_loc) Co_loc = Co;
        }
        s_vq[o] = vq; s_vn[o] = vn;
      }
      const Real Co_max = w.max_nonneg(Co_loc);
      // generalized_sedimentation: p3_upwind_impl.hpp:97-129
      if (!(Co_max >= 0)) err |= 4;
      const int tmpint1 = static_cast<int>(Co_max + 1);
      const Real dt_sub = P::div_by_count(dt_left, tmpint1);
      Real pl_bm = 0, pl_qm = 0, pl_n = 0, pl_q = 0;      // lane 0: flux out of the last lane of the previous slot
#pragma unroll 1
      for (int j = 0; j < nslot && kw + j * WS <= k_temp; ++j) {
        const int k = kw + lane + j * WS, o = j * WS;
        const bool in = k <= k_temp;
        const bool has_v = kw + j * WS <= k_qxbot;         // a slot the front enters in this sub-step has no fall speed yet
        const Real Vq = has_v ? s_vq[o] : Real(0), Vn = has_v ? s_vn[o] : Real(0);
        const Real rho = s_rho[o], inv_rho = s_irho[o], inv_dz = s_idz[o];
        Real r[4] = {s_bm[o], s_qm[o], s_n[o], s_q[o]};
        const Real V[4] = {Vq, Vq, Vn, Vq};
        Real pl[4] = {pl_bm, pl_qm, pl_n, pl_q};
        Real fq = 0;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const Real f = in ? V[i] * r[i] * rho : Real(0);
          // one rotation serves both: lanes > 0 receive the flux of the level above, lane 0 the flux out of this
          // slot's last lane, which is what lane 0 (and only lane 0) needs in the next slot
          const Real rot = w.rot_up1(f);
          const Real up = (lane == 0) ? pl[i] : rot;
          pl[i] = rot;
          const Real fluxdiv = (up - f) * inv_dz;
          const Real r_new = r[i] + fluxdiv * dt_sub * inv_rho;
          r[i] = in ? r_new : r[i];
          if (i == 3) fq = f;
        }
        pl_bm = pl[0]; pl_qm = pl[1]; pl_n = pl[2]; pl_q = pl[3];
        s_bm[o] = r[0]; s_qm[o] = r[1]; s_n[o] = r[2]; s_q[o] = r[3];
        const Real prt_new = prt_local + fq * dt_sub;
        prt_local = (k == kbot && k_qxbot == kbot) ? prt_new : prt_local;
      }
      if (k_qxbot != kbot) k_qxbot += 1;
      dt_left -= dt_sub;
      first = false;
      ++nsub;
    }
    // write back the part of the window the front reached
#pragma unroll 1
    for (int j = 0; j < nslot && kw + j * WS <= k_qxbot; ++j) {
      const int k = kw + lane + j * WS, o = j * WS;
      if (k <= k_qxbot) { g_q[k] = s_q[o]; g_n[k] = s_n[o]; g_qm[k] = s_qm[o]; g_bm[k] = s_bm[o]; }
    }
    const Real prt_accum = w.shfl_from(prt_local, (kbot - kw) % WS);   // only the lane owning kbot accumulated
    const Real surf = prt_accum * C::INV_RHO_H2O * inv_dt;
    err = w.or_int(err);
    if (lane == 0) {
      a.f[F_PRECIP_ICE_SURF][col] += surf;                // p3_ice_sed_impl.hpp:192 (only ice writes it)
      int* d = a.diag + col * D_NUM;
      d[D_QI_KTOP + 1] = k_qxbot; d[D_QI_KTOP + 2] = nsub;
      if (err) WP::atomic_or(d + D_ERROR, err);
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

  // a column that left p3_main after part1 or part2 (p3_main_impl.hpp:237,258) keeps the part3 defaults of
  // p3_main_init (p3_main_impl.hpp:40-47); this also overwrites the scratch value parked in diag_eff_radius_qc
  template <bool UNI>
  static P3_HD void defaults(const MainArgs<Real>& a, int64_t col, int k) {
    const ColAddr<UNI, Real> cp(a, col);
    cp.template p<F_DIAG_EFF_RADIUS_QC>(a)[k] = sp(10.e-6);
    cp.template p<F_DIAG_EFF_RADIUS_QI>(a)[k] = sp(25.e-6);
    cp.template p<F_DIAG_EFF_RADIUS_QR>(a)[k] = sp(500.e-6);
    cp.template p<F_RHO_QI>(a)[k] = 0;
  }

  // col_full: the column went through sedimentation (decided by the caller from the integer diagnostics; the
  // kernel passes the value of a load that is still in flight, so that it overlaps the eight loads below)
  template <bool UNI>
  static P3_HD void level(const MainArgs<Real>& a, int64_t col, int k, bool col_full = true) {
    const ColAddr<UNI, Real> cp(a, col);
    Real* const g_qc = cp.template p<F_QC>(a) + k; Real* const g_nc = cp.template p<F_NC>(a) + k;
    Real* const g_qr = cp.template p<F_QR>(a) + k; Real* const g_nr = cp.template p<F_NR>(a) + k;
    Real* const g_qi = cp.template p<F_QI>(a) + k; Real* const g_qm = cp.template p<F_QM>(a) + k;
    Real* const g_ni = cp.template p<F_NI>(a) + k; Real* const g_bm = cp.template p<F_BM>(a) + k;
    Real* const g_stash = cp.template p<F_DIAG_EFF_RADIUS_QC>(a) + k;
    const Real qc0 = *g_qc, nc0 = *g_nc, qr0 = *g_qr, nr0 = *g_nr, qi0 = *g_qi, qm0 = *g_qm, ni0 = *g_ni, bm0 = *g_bm;
    if (!col_full) { defaults<UNI>(a, col, k); return; }
    typename P::Part3Out o;
    o.eff_radius_qc = sp(10.e-6); o.eff_radius_qi = sp(25.e-6); o.eff_radius_qr = sp(500.e-6); o.rho_qi = 0;
    if (qc0 == 0 && qr0 == 0 && qi0 == 0) {
      // nothing to freeze, nothing to clip into vapour (qv + 0, th - x*0 are identities): part3 only
      // zeroes the number/rime fields of the absent species (p3_main_impl_part3.hpp:98-100,128-130,190-196,201)
      put(g_nc, Real(0), nc0); put(g_nr, Real(0), nr0); put(g_ni, Real(0), ni0);
      put(g_qm, Real(0), qm0); put(g_bm, Real(0), bm0);
    } else {
      Real* const g_qv = cp.template p<F_QV>(a) + k; Real* const g_th = cp.template p<F_TH>(a) + k;
      const Real qv0 = *g_qv, th0 = *g_th;
      Real qc = qc0, nc = nc0, qr = qr0, nr = nr0, qi = qi0, qm = qm0, ni = ni0, bm = bm0, qv = qv0, th = th0;
      const bool frz = *g_stash < 0;
      const Real inv_exner = P3_LDG(cp.template p<F_INV_EXNER>(a) + k);
      if (a.ro.do_ice_production)
        P::homogeneous_freezing_level(frz, inv_exner, qc, nc, qr, nr, qi, ni, qm, bm, th);
      // rho and the cloud fractions are only consumed where the species exists
      Real rho = 1, cf_l = 1, cf_r = 1, cf_i = 1;
      if (qc >= C::QSMALL || qi >= C::QSMALL)
        rho = P3_LDG(cp.template p<F_DPRES>(a) + k) / P3_LDG(cp.template p<F_DZ>(a) + k) / C::gravit;
      if (qc >= C::QSMALL) cf_l = P3_LDG(cp.template p<F_CLD_FRAC_L>(a) + k);
      if (qr >= C::QSMALL) cf_r = P3_LDG(cp.template p<F_CLD_FRAC_R>(a) + k);
      if (qi >= C::QSMALL) cf_i = P3_LDG(cp.template p<F_CLD_FRAC_I>(a) + k);
      Real sub_qc, sub_qr;
      P::part3_level(qc, nc, qr, nr, qi, ni, qm, bm, qv, th, rho, 1 / rho, inv_exner, cf_l, cf_r, cf_i, a.tb, a.ro, o, sub_qc, sub_qr);
      if (sub_qc != 0 || sub_qr != 0) {   // vap_liq_exchange read-modify-write, p3_main_impl_part3.hpp:97,127
        Real* const g_vle = cp.template p<F_VAP_LIQ_EXCHANGE>(a) + k;
        Real v = *g_vle;
        v = v - sub_qc;
        v = v - sub_qr;
        *g_vle = v;
      }
      put(g_qc, qc, qc0); put(g_nc, nc, nc0); put(g_qr, qr, qr0); put(g_nr, nr, nr0); put(g_qi, qi, qi0);
      put(g_qm, qm, qm0); put(g_ni, ni, ni0); put(g_bm, bm, bm0); put(g_qv, qv, qv0); put(g_th, th, th0);
    }
    *g_stash = o.eff_radius_qc;
    cp.template p<F_DIAG_EFF_RADIUS_QI>(a)[k] = o.eff_radius_qi;
    cp.template p<F_DIAG_EFF_RADIUS_QR>(a)[k] = o.eff_radius_qr;
    cp.template p<F_RHO_QI>(a)[k] = o.rho_qi;
  }
};

}  // namespace p3b200
#endif
