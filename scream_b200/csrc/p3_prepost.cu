// p3_prepost.cu — p3_preamble / p3_postamble of P3Microphysics::run_impl as sm_100a kernels, and the process's
// property checks folded into the postamble pass (p3_checks.cuh).
//
// Reference: the two functors of eamxx_p3_process_interface.hpp (p3_preamble :62-129, p3_postamble :200-262),
// run once per column with a serial loop over the level packs (Kokkos RangePolicy over columns,
// eamxx_p3_run.cpp:11-16, 39-44).  Here each is ONE streaming pass with one thread per (column, level):
// consecutive threads own consecutive levels of a column, so every access is a coalesced row segment; the only
// cross-level input (the max-overlap rain fraction needs cld_frac_t of the level above) is a second load of an
// address the neighbouring lane fetches anyway.  Both kernels are pure HBM streams:
//   preamble  15 reads + 16 writes per column-level (248 B in fp64)
//   postamble 17 reads + 15 writes                  (256 B in fp64)
// The helper formulas follow share/util/scream_common_physics_functions_impl.hpp (cited per line).
#include "p3_kernel.cuh"
#include "p3_prepost.cuh"
#include "../../include/p3_b200.h"

#include <cuda_runtime.h>
#include <cstdint>
#include <initializer_list>

#if !defined(P3B200_PP_MINB)
#define P3B200_PP_MINB 2     // resident 256-thread CTAs per SM the streaming kernels are compiled for
#endif

namespace p3b200 {

template <typename Real>
struct PreArgs {
  const Real *pmid, *pd, *pdd, *T_atm, *cld_frac_t;
  Real *q[10];                         // qv qc nc qr nr qi qm ni bm qv_prev
  Real *inv_exner, *th_atm, *cld_frac_l, *cld_frac_i, *cld_frac_r, *dz;
  int64_t stride;
  int ncol, nk;
};

// N consecutive levels per thread (N = 2: one 16-byte access per array and thread when levels, stride and base
// addresses are even), streaming cache hints: every line is touched exactly once.
template <typename Real, int N> struct PackIO;
template <typename Real> struct PackIO<Real, 1> {
  static __device__ __forceinline__ void ld(const Real* p, Real* v) { v[0] = __ldcs(p); }
  static __device__ __forceinline__ void st(Real* p, const Real* v) { __stcs(p, v[0]); }
};
template <> struct PackIO<double, 2> {
  static __device__ __forceinline__ void ld(const double* p, double* v) {
    const double2 t = __ldcs(reinterpret_cast<const double2*>(p)); v[0] = t.x; v[1] = t.y;
  }
  static __device__ __forceinline__ void st(double* p, const double* v) { __stcs(reinterpret_cast<double2*>(p), make_double2(v[0], v[1])); }
};
template <> struct PackIO<float, 2> {
  static __device__ __forceinline__ void ld(const float* p, float* v) {
    const float2 t = __ldcs(reinterpret_cast<const float2*>(p)); v[0] = t.x; v[1] = t.y;
  }
  static __device__ __forceinline__ void st(float* p, const float* v) { __stcs(reinterpret_cast<float2*>(p), make_float2(v[0], v[1])); }
};

template <typename Real, int N>
__global__ void __launch_bounds__(256, P3B200_PP_MINB)
p3_preamble_kernel(const __grid_constant__ PreArgs<Real> a) {
  using PP = PrePost<Real>;
  using C = Cst<Real>;
  using IO = PackIO<Real, N>;
  const int nkp = a.nk / N;                                  // packs per column (N divides nk)
  const int64_t total = (int64_t)a.ncol * nkp;
  for (int64_t idx = (int64_t)blockIdx.x * 256 + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * 256) {
    const int64_t col = idx / nkp;
    const int k = (int)(idx - col * nkp) * N;
    const int64_t o = col * a.stride + k;
    Real pmid[N], pd[N], pdd[N], T[N], cft[N], q[10][N];
    IO::ld(a.pmid + o, pmid); IO::ld(a.pd + o, pd); IO::ld(a.pdd + o, pdd); IO::ld(a.T_atm + o, T);
    IO::ld(a.cld_frac_t + o, cft);
    const Real cft_above = (k != 0) ? __ldg(a.cld_frac_t + o - 1) : Real(0);
#pragma unroll
    for (int i = 0; i < 10; ++i) IO::ld(a.q[i] + o, q[i]);
    Real dz[N], inv_ex[N], th[N], cf[N], cfr[N];
#pragma unroll
    for (int j = 0; j < N; ++j) {
      // calculate_dz :400-409 with calculate_virtual_temperature :214-222, on the WET qv (preamble l.78)
      const Real c1 = -Real(1) + Real(1) / C::ep_2;
      const Real T_virtual = T[j] * (Real(1) + c1 * q[0][j]);
      dz[j] = (C::RD / C::gravit) * pd[j] * T_virtual / pmid[j];
      // wet -> dry (l.89-101)
#pragma unroll
      for (int i = 0; i < 10; ++i) q[i][j] = PP::dry(q[i][j], pd[j], pdd[j]);
      // exner, theta (l.104-107; calculate_theta_from_T :115-118)
      const Real ex = PP::exner(pmid[j]);
      inv_ex[j] = Real(1.0) / ex;
      th[j] = T[j] / ex;
      // cloud fractions (l.110-112) and max-overlap rain fraction (l.115-129)
      const Real mincld = Real(0.0001);
      cf[j] = cft[j] < mincld ? mincld : cft[j];
      const Real up = j == 0 ? cft_above : cft[j == 0 ? 0 : j - 1];
      cfr[j] = (k + j != 0 && up > cf[j]) ? up : cf[j];
    }
    IO::st(a.dz + o, dz);
#pragma unroll
    for (int i = 0; i < 10; ++i) IO::st(a.q[i] + o, q[i]);
    IO::st(a.inv_exner + o, inv_ex); IO::st(a.th_atm + o, th);
    IO::st(a.cld_frac_l + o, cf); IO::st(a.cld_frac_i + o, cf); IO::st(a.cld_frac_r + o, cfr);
  }
}

// One thread per (column, level); the body is postamble_level (p3_prepost.cuh), shared with the fused tail of phase C.
template <typename Real, bool CHECK>
__global__ void __launch_bounds__(256)
p3_postamble_kernel(const __grid_constant__ PostArgs<Real> a) {
  CheckFlags<Real> chk;
  chk.init();
  const int64_t total = (int64_t)a.ncol * a.nk;
  for (int64_t idx = (int64_t)blockIdx.x * 256 + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * 256) {
    const int64_t col = idx / a.nk;
    const int k = (int)(idx - col * a.nk);
    postamble_level<Real, CHECK, false>(a, col, k, chk);
  }
  if (CHECK) chk.publish(a.check_mask);
}

// The same statements on two consecutive levels per thread with 16-byte streaming accesses (see PackIO).
template <typename Real, bool CHECK>
__global__ void __launch_bounds__(256, CHECK ? P3B200_PP_MINB : P3B200_PP_MINB + 1)   // measured 2/3/4 CTAs per SM: plain 3.33/3.05/3.49 ms, with the check mask 3.51/3.54/5.10 ms
p3_postamble_kernel2(const __grid_constant__ PostArgs<Real> a) {
  using PP = PrePost<Real>;
  using C = Cst<Real>;
  using IO = PackIO<Real, 2>;
  CheckFlags<Real> chk;
  chk.init();
  const int nkp = a.nk / 2;
  const int64_t total = (int64_t)a.ncol * nkp;
  for (int64_t idx = (int64_t)blockIdx.x * 256 + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * 256) {
    const int64_t col = idx / nkp;
    const int k = (int)(idx - col * nkp) * 2;
    const int64_t o = col * a.stride + k;
    Real pmid[2], pd[2], pdd[2], th[2], T[2], w[9][2], r[3][2];
    IO::ld(a.pmid + o, pmid); IO::ld(a.pd + o, pd); IO::ld(a.pdd + o, pdd); IO::ld(a.th_atm + o, th);
    IO::ld(a.T_atm + o, T);
#pragma unroll
    for (int i = 0; i < 9; ++i) IO::ld(a.q[i] + o, w[i]);
#pragma unroll
    for (int i = 0; i < 3; ++i) IO::ld(a.rad[i] + o, r[i]);
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const Real T_before = T[j];
      Real Tn = (th[j] * PP::exner(pmid[j]) - T_before) * pdd[j] / pd[j];
      Tn += T_before;
      T[j] = Tn;
#pragma unroll
      for (int i = 0; i < 9; ++i) w[i][j] = PP::wet(w[i][j], pd[j], pdd[j]);
#pragma unroll
      for (int i = 0; i < 3; ++i) r[i][j] = r[i][j] * Real(1e6);
      if (CHECK) {
        chk.template see<CHK_T_MID>(Tn);
        chk.template see<CHK_QC>(w[0][j]); chk.template see<CHK_NC>(w[1][j]); chk.template see<CHK_QR>(w[2][j]);
        chk.template see<CHK_NR>(w[3][j]); chk.template see<CHK_QI>(w[4][j]); chk.template see<CHK_NI>(w[5][j]);
        chk.template see<CHK_QM>(w[6][j]); chk.template see<CHK_BM>(w[7][j]); chk.template see<CHK_QV>(w[8][j]);
        chk.template see<CHK_RAD_QC>(r[0][j]); chk.template see<CHK_RAD_QI>(r[1][j]); chk.template see<CHK_RAD_QR>(r[2][j]);
      }
    }
    IO::st(a.T_atm + o, T); IO::st(a.T_prev + o, T);
#pragma unroll
    for (int i = 0; i < 9; ++i) IO::st(a.q[i] + o, w[i]);
    IO::st(a.qv_prev + o, w[8]);
#pragma unroll
    for (int i = 0; i < 3; ++i) IO::st(a.rad[i] + o, r[i]);
    if (k == 0) {
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
  if (CHECK) chk.publish(a.check_mask);
}

// ---- the property checks on their own (for a caller that does not use the fused postamble) ----------------
template <typename Real>
struct CheckArgs {
  const Real* f[CHK_NUM];              // CheckId order; CHK_PRECIP_* are per-column arrays
  int64_t stride;
  int ncol, nk;
  unsigned long long* state;
};

// One launch for all 15 reductions: blockIdx.y picks the check, the blocks of a row stream that field with four
// loads in flight per thread and keep ONE (min, max) pair.
template <typename Real>
__global__ void __launch_bounds__(256)
p3_checks_kernel(const __grid_constant__ CheckArgs<Real> a) {
  const int fid = blockIdx.y;
  const bool column_field = fid == CHK_PRECIP_LIQ || fid == CHK_PRECIP_ICE;
  const Real* __restrict__ f = a.f[fid];
  const int nk = column_field ? 1 : a.nk;
  const int64_t stride = column_field ? 1 : a.stride;
  const int64_t total = (int64_t)a.ncol * nk;
  const int64_t step = (int64_t)gridDim.x * 256;
  const bool contiguous = stride == nk;
  CheckMinMax<Real> mm;
  mm.init();
  if (contiguous && (total & 1) == 0 && reinterpret_cast<uintptr_t>(f) % (2 * sizeof(Real)) == 0) {
    // dense field: 16-byte streaming loads, four in flight per thread
    const int64_t pairs = total >> 1;
    for (int64_t base = (int64_t)blockIdx.x * 256 + threadIdx.x; base < pairs; base += 4 * step) {
      Real v[4][2];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int64_t idx = base + j * step;
        v[j][0] = v[j][1] = CheckMinMax<Real>::skipped();
        if (idx < pairs) PackIO<Real, 2>::ld(f + 2 * idx, v[j]);
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) { mm.see(v[j][0]); mm.see(v[j][1]); }
    }
  } else {
    for (int64_t base = (int64_t)blockIdx.x * 256 + threadIdx.x; base < total; base += 4 * step) {
      Real v[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int64_t idx = base + j * step;
        v[j] = CheckMinMax<Real>::skipped();
        if (idx < total) {
          const int64_t o = contiguous ? idx : (idx / nk) * stride + idx % nk;
          v[j] = __ldg(f + o);
        }
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) mm.see(v[j]);
    }
  }
  mm.template publish<256>(a.state + 2 * fid);
}

// state[2i] = key(+big), state[2i+1] = key(-big): the Kokkos MinMaxLoc identities
__global__ void p3_checks_reset_kernel(unsigned long long* state, double big) {
  if (threadIdx.x < 2 * CHK_NUM) state[threadIdx.x] = chk_encode((threadIdx.x & 1) ? -big : big);
}

// Lowest flat index (col * nk + k, the reference's unflattened (col, lev) order) whose value equals `target`.
// Only launched for a field whose check did not pass.
template <typename Real>
__global__ void __launch_bounds__(256)
p3_find_loc_kernel(const Real* __restrict__ f, int64_t stride, int ncol, int nk, double target, unsigned long long* loc) {
  const int64_t total = (int64_t)ncol * nk;
  unsigned long long best = ~0ull;
  for (int64_t idx = (int64_t)blockIdx.x * 256 + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * 256) {
    const int64_t col = idx / nk;
    const int k = (int)(idx - col * nk);
    if ((double)__ldg(f + col * stride + k) == target && (unsigned long long)idx < best) best = (unsigned long long)idx;
  }
  if (best != ~0ull) atomicMin(loc, best);
}

// FieldWithinIntervalCheck::repair_impl (field_within_interval_check.cpp:338-372): clip into [lb, ub]
template <typename Real>
__global__ void __launch_bounds__(256)
p3_clip_kernel(Real* __restrict__ f, int64_t stride, int ncol, int nk, Real lb, Real ub) {
  const int64_t total = (int64_t)ncol * nk;
  for (int64_t idx = (int64_t)blockIdx.x * 256 + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * 256) {
    const int64_t col = idx / nk;
    const int k = (int)(idx - col * nk);
    Real v = f[col * stride + k];
    v = ub < v ? ub : v;
    v = lb > v ? lb : v;
    f[col * stride + k] = v;
  }
}

static int grid_for(long long total) {
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  long long blocks = (total + 255) / 256;
  const long long cap = (long long)sms * 8 * 4;      // a few waves of 8 resident CTAs per SM
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return (int)blocks;
}

// 16-byte accesses need even levels/stride and 16-byte aligned bases
template <typename Real>
static bool pack2_ok(int nk, int64_t stride, std::initializer_list<const void*> ptrs) {
  if ((nk & 1) || (stride & 1)) return false;
  for (const void* p : ptrs) if (reinterpret_cast<uintptr_t>(p) % (2 * sizeof(Real))) return false;
  return true;
}

template <typename Real>
int launch_preamble(const PreArgs<Real>& a, void* stream) {
  if (a.ncol <= 0) return 0;
  const bool two = pack2_ok<Real>(a.nk, a.stride, {a.pmid, a.pd, a.pdd, a.T_atm, a.cld_frac_t, a.q[0], a.q[1], a.q[2], a.q[3],
                                                   a.q[4], a.q[5], a.q[6], a.q[7], a.q[8], a.q[9], a.inv_exner, a.th_atm,
                                                   a.cld_frac_l, a.cld_frac_i, a.cld_frac_r, a.dz});
  if (two) p3_preamble_kernel<Real, 2><<<grid_for((long long)a.ncol * (a.nk / 2)), 256, 0, (cudaStream_t)stream>>>(a);
  else p3_preamble_kernel<Real, 1><<<grid_for((long long)a.ncol * a.nk), 256, 0, (cudaStream_t)stream>>>(a);
  return (int)cudaGetLastError();
}
template <typename Real>
int launch_postamble(const PostArgs<Real>& a, void* stream) {
  if (a.ncol <= 0) return 0;
  const bool two = pack2_ok<Real>(a.nk, a.stride, {a.pmid, a.pd, a.pdd, a.th_atm, a.T_atm, a.T_prev, a.qv_prev, a.q[0], a.q[1],
                                                   a.q[2], a.q[3], a.q[4], a.q[5], a.q[6], a.q[7], a.q[8], a.rad[0], a.rad[1],
                                                   a.rad[2]});
  const int grid = grid_for((long long)a.ncol * (two ? a.nk / 2 : a.nk));
  cudaStream_t st = (cudaStream_t)stream;
  if (two) {
    if (a.check_mask) p3_postamble_kernel2<Real, true><<<grid, 256, 0, st>>>(a);
    else p3_postamble_kernel2<Real, false><<<grid, 256, 0, st>>>(a);
  } else {
    if (a.check_mask) p3_postamble_kernel<Real, true><<<grid, 256, 0, st>>>(a);
    else p3_postamble_kernel<Real, false><<<grid, 256, 0, st>>>(a);
  }
  return (int)cudaGetLastError();
}

template <typename Real>
int preamble_from_abi(const P3B200PreambleFields* f, int ncol, int nlev, void* stream) {
  PreArgs<Real> a;
  a.pmid = (const Real*)f->pmid; a.pd = (const Real*)f->pseudo_density; a.pdd = (const Real*)f->pseudo_density_dry;
  a.T_atm = (const Real*)f->T_atm; a.cld_frac_t = (const Real*)f->cld_frac_t;
  void* q[10] = {f->qv, f->qc, f->nc, f->qr, f->nr, f->qi, f->qm, f->ni, f->bm, f->qv_prev};
  for (int i = 0; i < 10; ++i) a.q[i] = (Real*)q[i];
  a.inv_exner = (Real*)f->inv_exner; a.th_atm = (Real*)f->th_atm; a.cld_frac_l = (Real*)f->cld_frac_l;
  a.cld_frac_i = (Real*)f->cld_frac_i; a.cld_frac_r = (Real*)f->cld_frac_r; a.dz = (Real*)f->dz;
  a.stride = f->col_stride; a.ncol = ncol; a.nk = nlev;
  return launch_preamble<Real>(a, stream);
}
template <typename Real>
PostArgs<Real> make_post_args(const P3B200PostambleFields* f, int ncol, int nlev, unsigned* check_mask) {
  PostArgs<Real> a;
  a.check_mask = check_mask;
  a.pmid = (const Real*)f->pmid; a.pd = (const Real*)f->pseudo_density; a.pdd = (const Real*)f->pseudo_density_dry;
  a.th_atm = (const Real*)f->th_atm; a.T_atm = (Real*)f->T_atm; a.T_prev = (Real*)f->T_prev; a.qv_prev = (Real*)f->qv_prev;
  void* q[9] = {f->qc, f->nc, f->qr, f->nr, f->qi, f->ni, f->qm, f->bm, f->qv};
  for (int i = 0; i < 9; ++i) a.q[i] = (Real*)q[i];
  a.rad[0] = (Real*)f->diag_eff_radius_qc; a.rad[1] = (Real*)f->diag_eff_radius_qi; a.rad[2] = (Real*)f->diag_eff_radius_qr;
  a.liq_flux = (const Real*)f->precip_liq_surf_flux; a.ice_flux_in = (const Real*)f->precip_ice_surf_flux;
  a.liq_mass = (Real*)f->precip_liq_surf_mass; a.ice_mass = (Real*)f->precip_ice_surf_mass;
  a.vapor_flux = (Real*)f->vapor_flux; a.water_flux = (Real*)f->water_flux; a.ice_flux = (Real*)f->ice_flux;
  a.heat_flux = (Real*)f->heat_flux;
  a.stride = f->col_stride; a.ncol = ncol; a.nk = nlev; a.dt = f->dt;
  return a;
}
template <typename Real>
int postamble_from_abi(const P3B200PostambleFields* f, int ncol, int nlev, void* stream, unsigned* check_mask) {
  return launch_postamble<Real>(make_post_args<Real>(f, ncol, nlev, check_mask), stream);
}

int checks_reset(unsigned long long* state, int real_bytes, void* stream) {
  p3_checks_reset_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(state, real_bytes == 8 ? DBL_MAX : (double)FLT_MAX);
  return (int)cudaGetLastError();
}

template <typename Real>
int checks_from_abi(const P3B200CheckFields* f, int ncol, int nlev, void* stream, unsigned long long* state) {
  if (ncol <= 0) return 0;
  CheckArgs<Real> a;
  const void* src[CHK_NUM] = {f->T_mid, f->qv, f->qc, f->qi, f->qr, f->qm, f->nc, f->nr, f->ni, f->bm,
                              f->precip_liq_surf_mass, f->precip_ice_surf_mass, f->eff_radius_qc, f->eff_radius_qi,
                              f->eff_radius_qr};
  for (int i = 0; i < CHK_NUM; ++i) a.f[i] = (const Real*)src[i];
  a.stride = f->col_stride; a.ncol = ncol; a.nk = nlev; a.state = state;
  int gx = grid_for(((long long)ncol * nlev + 7) / 8);
  if (gx > 148 * 8) gx = 148 * 8;       // x CHK_NUM rows: a few waves of 8 resident CTAs per SM
  p3_checks_kernel<Real><<<dim3(gx, CHK_NUM), 256, 0, (cudaStream_t)stream>>>(a);
  return (int)cudaGetLastError();
}

template <typename Real>
int find_loc(const void* field, int64_t stride, int ncol, int nk, double target, unsigned long long* loc_dev, void* stream) {
  p3_find_loc_kernel<Real><<<grid_for((long long)ncol * nk), 256, 0, (cudaStream_t)stream>>>(
      (const Real*)field, stride, ncol, nk, target, loc_dev);
  return (int)cudaGetLastError();
}

template <typename Real>
int clip_field(void* field, int64_t stride, int ncol, int nk, double lb, double ub, void* stream) {
  if (ncol <= 0) return 0;
  p3_clip_kernel<Real><<<grid_for((long long)ncol * nk), 256, 0, (cudaStream_t)stream>>>(
      (Real*)field, stride, ncol, nk, (Real)lb, (Real)ub);
  return (int)cudaGetLastError();
}

template int launch_postamble<double>(const PostArgs<double>&, void*);
template int launch_postamble<float>(const PostArgs<float>&, void*);
template int preamble_from_abi<double>(const P3B200PreambleFields*, int, int, void*);
template int preamble_from_abi<float>(const P3B200PreambleFields*, int, int, void*);
template PostArgs<double> make_post_args<double>(const P3B200PostambleFields*, int, int, unsigned*);
template PostArgs<float> make_post_args<float>(const P3B200PostambleFields*, int, int, unsigned*);
template int postamble_from_abi<double>(const P3B200PostambleFields*, int, int, void*, unsigned*);
template int postamble_from_abi<float>(const P3B200PostambleFields*, int, int, void*, unsigned*);
template int checks_from_abi<double>(const P3B200CheckFields*, int, int, void*, unsigned long long*);
template int checks_from_abi<float>(const P3B200CheckFields*, int, int, void*, unsigned long long*);
template int find_loc<double>(const void*, int64_t, int, int, double, unsigned long long*, void*);
template int find_loc<float>(const void*, int64_t, int, int, double, unsigned long long*, void*);
template int clip_field<double>(void*, int64_t, int, int, double, double, void*);
template int clip_field<float>(void*, int64_t, int, int, double, double, void*);

}  // namespace p3b200
