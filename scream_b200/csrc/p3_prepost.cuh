// p3_prepost.cuh — pieces of the p3_postamble shared by the stand-alone kernel (p3_prepost.cu) and the fused
// tail of phase C (p3_kernel.cu): argument block, helper formulas, the per-level body.
//
// Reference: p3_postamble, eamxx_p3_process_interface.hpp:200-262 (helper formulas:
// share/util/scream_common_physics_functions_impl.hpp, cited per line).
#pragma once
#include "p3_physics.cuh"
#include "p3_checks.cuh"

#include <cstdint>

struct P3B200PostambleFields;

namespace p3b200 {

template <typename Real>
struct PostArgs {
  const Real *pmid, *pd, *pdd, *th_atm;
  Real *T_atm, *T_prev, *qv_prev;
  Real *q[9];                          // qc nc qr nr qi ni qm bm qv  (order of the reference, l.225-233)
  Real *rad[3];
  const Real *liq_flux, *ice_flux_in;
  Real *liq_mass, *ice_mass, *vapor_flux, *water_flux, *ice_flux, *heat_flux;
  int64_t stride;
  int ncol, nk;
  double dt;
  unsigned* check_mask;                // CHECK only: the violation mask word (p3_checks.cuh)
};

// C-ABI struct -> typed argument block (p3_prepost.cu)
template <typename Real>
PostArgs<Real> make_post_args(const P3B200PostambleFields* f, int ncol, int nlev, unsigned* check_mask);

// p3_postamble over all levels of `a` on `stream` (p3_prepost.cu); folds the property-check mask in when a.check_mask
template <typename Real>
int launch_postamble(const PostArgs<Real>& a, void* stream);

#if defined(__CUDACC__)
template <typename Real>
struct PrePost {
  using C = Cst<Real>;
  // exner_function :88-97
  static __device__ __forceinline__ Real exner(Real p) { return Phys<Real>::x_pow(p / C::P0, C::RD * C::INV_CP); }
  // calculate_drymmr_from_wetmmr_dp_based :375-379 / calculate_wetmmr_from_drymmr_dp_based :327-331
  // div_z: most mixing ratios of most levels are exactly 0, and a zero numerator sends the fp64 division out of line
  static __device__ __forceinline__ Real dry(Real wet, Real pd, Real pdd) { return Phys<Real>::div_z(wet * pd, pdd); }
  static __device__ __forceinline__ Real wet(Real dry, Real pd, Real pdd) { return Phys<Real>::div_z(dry * pdd, pd); }
};


// The postamble of one (column, level).  CHECK folds the 15 property checks of initialize_impl (:226-243) into
// the pass: they are evaluated on the values just produced (what the reference's post-condition / invariant checks
// read from memory afterwards).  FUSED = called at the end of phase C by the thread that has just written th, the
// mixing ratios and the radii of this level: those are re-read with plain loads (never through the read-only path).
template <typename Real, bool CHECK, bool FUSED>
__device__ __forceinline__ void postamble_level(const PostArgs<Real>& a, int64_t col, int k, CheckFlags<Real>& chk) {
  using PP = PrePost<Real>;
  using C = Cst<Real>;
  const int64_t o = col * a.stride + k;
  const Real pmid = __ldg(a.pmid + o), pd = __ldg(a.pd + o), pdd = __ldg(a.pdd + o);
  // temperature update with the dry/wet rescaling of dT (l.208-216; calculate_T_from_theta :162-165)
  const Real T_before = a.T_atm[o];
  const Real th = FUSED ? a.th_atm[o] : __ldg(a.th_atm + o);
  Real T = (th * PP::exner(pmid) - T_before) * pdd / pd;
  T += T_before;
  a.T_atm[o] = T;
  a.T_prev[o] = T;
  if (CHECK) chk.template see<CHK_T_MID>(T);
  // dry -> wet (l.225-234)
  Real w[9];
#pragma unroll
  for (int i = 0; i < 9; ++i) {
    w[i] = PP::wet(a.q[i][o], pd, pdd);
    a.q[i][o] = w[i];
  }
  a.qv_prev[o] = w[8];
  // effective radii in microns (l.237-239)
  Real r[3];
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    r[i] = a.rad[i][o] * Real(1e6);
    a.rad[i][o] = r[i];
  }
  if (CHECK) {   // q[] order: qc nc qr nr qi ni qm bm qv
    chk.template see<CHK_QC>(w[0]); chk.template see<CHK_NC>(w[1]); chk.template see<CHK_QR>(w[2]);
    chk.template see<CHK_NR>(w[3]); chk.template see<CHK_QI>(w[4]); chk.template see<CHK_NI>(w[5]);
    chk.template see<CHK_QM>(w[6]); chk.template see<CHK_BM>(w[7]); chk.template see<CHK_QV>(w[8]);
    chk.template see<CHK_RAD_QC>(r[0]); chk.template see<CHK_RAD_QI>(r[1]); chk.template see<CHK_RAD_QR>(r[2]);
  }
  if (k == 0) {
    // accumulated surface masses and boundary fluxes (l.244-258); m_dt is a double in the reference
    const Real lf = a.liq_flux[col], icf = a.ice_flux_in[col];
    const Real lm = Real(a.liq_mass[col] + lf * C::RHO_H2O * a.dt);
    const Real im = Real(a.ice_mass[col] + icf * C::RHO_H2O * a.dt);
    a.liq_mass[col] = lm;
    a.ice_mass[col] = im;
    if (CHECK) { chk.template see<CHK_PRECIP_LIQ>(lm); chk.template see<CHK_PRECIP_ICE>(im); }
    if (a.water_flux) {
      a.vapor_flux[col] = 0;
      a.water_flux[col] = lf + icf;
      a.ice_flux[col] = icf;
      a.heat_flux[col] = 0;
    }
  }
}
#endif  // __CUDACC__

}  // namespace p3b200
