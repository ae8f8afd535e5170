// p3_budget.cu — column and domain totals of water mass and moist energy for the global conservation
// check that brackets p3_main.
//
// Per column (mass_and_energy_column_conservation_check.cpp:283-363, with the terms P3 cannot change —
// kinetic energy and phis*ps/g — left out because they cancel in the before/after difference):
//   water  = sum_k (qv+qc+qi+qr) * pseudo_density / g                       (:295-301)
//   energy = sum_k (T*Cpair + (LatVap+LatIce)*qv + LatIce*(qc+qr)) * pseudo_density / g   (:335-346), T = th/inv_exner
//   water that left through the surface  = RHO_H2O * (precip_liq_surf + precip_ice_surf) * dt    (:312)
//   energy that left through the surface = RHO_H2O * LatIce * precip_liq_surf * dt               (:362)
// All sums are formed in double in a FIXED order (lane-strided partial sums, xor butterfly, then a fixed
// two-level tree over columns), so a given shard always produces the same bits; the cross-rank sum is one
// all-reduce of 4 doubles done by the caller (NCCL).
#include "p3_kernel.cuh"

#include <cuda_runtime.h>

#include <mutex>

namespace p3b200 {

constexpr int BUD_WARPS = 8;
constexpr int BUD_THREADS = 256;

template <typename Real>
__global__ void __launch_bounds__(BUD_WARPS * 32)
p3_column_budget(const __grid_constant__ MainArgs<Real> a, double* __restrict__ col_vals /* [ncol][4] */) {
  const int lane = threadIdx.x & 31;
  const int64_t col = (int64_t)blockIdx.x * BUD_WARPS + (threadIdx.x >> 5);
  if (col >= a.ncol) return;
  using C = Cst<double>;
  const Real* qv = col_ptr(a, F_QV, col); const Real* qc = col_ptr(a, F_QC, col);
  const Real* qi = col_ptr(a, F_QI, col); const Real* qr = col_ptr(a, F_QR, col);
  const Real* th = col_ptr(a, F_TH, col); const Real* ie = col_ptr(a, F_INV_EXNER, col);
  const Real* dp = col_ptr(a, F_DPRES, col);
  double w = 0, e = 0;
  for (int k = lane; k < a.nk; k += 32) {
    const double m = (double)dp[k] / C::gravit;
    const double v = qv[k], c = qc[k], i = qi[k], r = qr[k];
    const double T = (double)th[k] / (double)ie[k];
    w += (v + c + i + r) * m;
    e += (T * C::Cpair + (C::LatVap + C::LatIce) * v + C::LatIce * (c + r)) * m;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    w += __shfl_xor_sync(0xffffffffu, w, o);
    e += __shfl_xor_sync(0xffffffffu, e, o);
  }
  if (lane == 0) {
    const double pls = a.f[F_PRECIP_LIQ_SURF][col], pis = a.f[F_PRECIP_ICE_SURF][col];
    double* o = col_vals + col * 4;
    o[0] = w; o[1] = e;
    o[2] = C::RHO_H2O * (pls + pis) * (double)a.dt;
    o[3] = C::RHO_H2O * C::LatIce * pls * (double)a.dt;
  }
}

// fixed-order sum of rows [begin, end) of vals[n][4]: thread t adds rows begin+t, begin+t+T, ... then a
// shared-memory tree with a fixed shape
__device__ __forceinline__ void block_sum4(const double* __restrict__ vals, int64_t begin, int64_t end, double* out4) {
  __shared__ double sm[BUD_THREADS][4];
  double s[4] = {0, 0, 0, 0};
  for (int64_t i = begin + threadIdx.x; i < end; i += BUD_THREADS)
    for (int j = 0; j < 4; ++j) s[j] += vals[i * 4 + j];
  for (int j = 0; j < 4; ++j) sm[threadIdx.x][j] = s[j];
  __syncthreads();
  for (int h = BUD_THREADS / 2; h > 0; h >>= 1) {
    if ((int)threadIdx.x < h)
      for (int j = 0; j < 4; ++j) sm[threadIdx.x][j] += sm[threadIdx.x + h][j];
    __syncthreads();
  }
  if (threadIdx.x == 0)
    for (int j = 0; j < 4; ++j) out4[j] = sm[0][j];
}

__global__ void __launch_bounds__(BUD_THREADS)
p3_budget_partials(const double* __restrict__ col_vals, int64_t ncol, int64_t per_block, double* __restrict__ partials) {
  const int64_t b = (int64_t)blockIdx.x * per_block;
  const int64_t e = b + per_block < ncol ? b + per_block : ncol;
  block_sum4(col_vals, b, e, partials + (int64_t)blockIdx.x * 4);
}

__global__ void __launch_bounds__(BUD_THREADS)
p3_budget_final(const double* __restrict__ partials, int nparts, double* __restrict__ totals4) {
  block_sum4(partials, 0, nparts, totals4);
}

template <typename Real>
int launch_budget(const MainArgs<Real>& a, double* col_vals, double* partials, int nparts, double* totals4, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (a.ncol <= 0) return (int)cudaMemsetAsync(totals4, 0, 4 * sizeof(double), stream);
  const int blocks = (a.ncol + BUD_WARPS - 1) / BUD_WARPS;
  p3_column_budget<Real><<<blocks, BUD_WARPS * 32, 0, stream>>>(a, col_vals);
  const int64_t per_block = ((int64_t)a.ncol + nparts - 1) / nparts;
  p3_budget_partials<<<nparts, BUD_THREADS, 0, stream>>>(col_vals, a.ncol, per_block, partials);
  p3_budget_final<<<1, BUD_THREADS, 0, stream>>>(partials, nparts, totals4);
  return (int)cudaGetLastError();
}

// =================================================================================================
// MassAndEnergyColumnConservationCheck (share/property_checks/mass_and_energy_column_conservation_check.cpp): the
// atmosphere driver calls compute_current_mass / compute_current_energy before a process runs (:56-107) and check()
// after it (:110-215).  Per column, with the within-column sums formed level by level in index order (the reference
// serialises them under ekatBFB, ekat_kokkos_utils.hpp:248):
//   mass   = sum_k (qv+qc+qi+qr) * pseudo_density / g                                                    (:283-301)
//   energy = sum_k (T*Cpair + 0.5*(u^2+v^2) + (LatVap+LatIce)*qv + LatIce*(qc+qr)) * pseudo_density / g + phis*ps/g   (:314-347)
//   check: rel_err_mass   = |mass - (mass_prev + (vapor_flux - water_flux*RHO_H2O)*dt)| / mass_prev       (:157-170)
//          rel_err_energy = |energy - (energy_prev + (vapor_flux*(LatVap+LatIce) - (water_flux-ice_flux)*RHO_H2O*LatIce
//                                                    + heat_flux)*dt)| / energy_prev                      (:185-199)
//   and the MaxLoc of each over the columns (`rel_err > result.val`: NaN never wins; ties: lowest column).
// One warp per column: the 32 lanes evaluate 32 levels' terms from coalesced loads, lane 0 adds them in level order
// (one shuffle per level), so the result has the bits of the sequential sum.
template <typename Real, bool CHECK>
__global__ void __launch_bounds__(BUD_WARPS * 32)
p3_column_conservation(const ConsArgs<Real> a, double* __restrict__ mass, double* __restrict__ energy,
                       double* __restrict__ rel_mass, double* __restrict__ rel_energy) {
  const int lane = threadIdx.x & 31;
  const int64_t col = (int64_t)blockIdx.x * BUD_WARPS + (threadIdx.x >> 5);
  if (col >= a.ncol) return;
  using C = Cst<Real>;
  const int64_t o = col * a.col_stride;
  const Real* u = a.horiz_winds + col * a.wind_col_stride;
  const Real* v = u + a.wind_comp_stride;
  Real tm = 0, te = 0;
  for (int k0 = 0; k0 < a.nk; k0 += 32) {
    const int k = k0 + lane;
    Real m_term = 0, e_term = 0;
    if (k < a.nk) {
      const Real qv = a.qv[o + k], qc = a.qc[o + k], qi = a.qi[o + k], qr = a.qr[o + k], pd = a.pseudo_density[o + k];
      m_term = (qv + qc + qi + qr) * pd / C::gravit;
      const Real u2 = u[k] * u[k], v2 = v[k] * v[k];
      e_term = (a.T_mid[o + k] * C::Cpair + Real(0.5) * (u2 + v2) + (C::LatVap + C::LatIce) * qv + C::LatIce * (qc + qr)) * pd / C::gravit;
    }
    const int n = (a.nk - k0 < 32) ? a.nk - k0 : 32;
    for (int j = 0; j < n; ++j) {
      tm += __shfl_sync(0xffffffffu, m_term, j);
      te += __shfl_sync(0xffffffffu, e_term, j);
    }
  }
  if (lane != 0) return;
  te += a.phis[col] * a.ps[col] / C::gravit;
  if (!CHECK) { mass[col] = (double)tm; energy[col] = (double)te; return; }
  const Real dt = a.dt;
  const Real prev_m = (Real)mass[col], prev_e = (Real)energy[col];
  const Real vf = a.vapor_flux[col], wf = a.water_flux[col], icf = a.ice_flux[col], hf = a.heat_flux[col];
  const Real tm_exp = prev_m + (vf - wf * C::RHO_H2O) * dt;
  const Real te_exp = prev_e + (vf * (C::LatVap + C::LatIce) - (wf - icf) * C::RHO_H2O * C::LatIce + hf) * dt;
  rel_mass[col] = (double)(std::abs(tm - tm_exp) / prev_m);
  rel_energy[col] = (double)(std::abs(te - te_exp) / prev_e);
}

// MaxLoc over the columns for both error arrays: out[0] = max mass error, out[1] = its column, out[2], out[3] energy.
// Starts from the reduction identity (lowest finite value, column -1); `v > best` as in the reference.
__global__ void __launch_bounds__(1024)
p3_conservation_maxloc(const double* __restrict__ rel_mass, const double* __restrict__ rel_energy, int64_t ncol,
                       double* __restrict__ out4, int real_bytes) {
  __shared__ double sv[2][1024];
  __shared__ long long si[2][1024];
  const double ident = real_bytes == 4 ? -(double)3.402823466e+38f : -1.7976931348623157e+308;
  double best[2] = {ident, ident};
  long long loc[2] = {-1, -1};
  for (int64_t i = threadIdx.x; i < ncol; i += 1024) {
    const double m = rel_mass[i], e = rel_energy[i];
    if (m > best[0]) { best[0] = m; loc[0] = i; }
    if (e > best[1]) { best[1] = e; loc[1] = i; }
  }
  for (int w = 0; w < 2; ++w) { sv[w][threadIdx.x] = best[w]; si[w][threadIdx.x] = loc[w]; }
  __syncthreads();
  for (int h = 512; h > 0; h >>= 1) {
    if ((int)threadIdx.x < h)
      for (int w = 0; w < 2; ++w) {
        const double ov = sv[w][threadIdx.x + h];
        const long long oi = si[w][threadIdx.x + h];
        const double mv = sv[w][threadIdx.x];
        const long long mi = si[w][threadIdx.x];
        // larger value wins; on equal values the lower column (what a serial sweep reports)
        if (ov > mv || (ov == mv && oi >= 0 && (mi < 0 || oi < mi))) { sv[w][threadIdx.x] = ov; si[w][threadIdx.x] = oi; }
      }
    __syncthreads();
  }
  if (threadIdx.x == 0) { out4[0] = sv[0][0]; out4[1] = (double)si[0][0]; out4[2] = sv[1][0]; out4[3] = (double)si[1][0]; }
}

template <typename Real>
int launch_conservation(const ConsArgs<Real>& a, bool check, double* mass, double* energy, double* rel_mass,
                        double* rel_energy, double* out4, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  if (a.ncol <= 0) return 0;
  const int blocks = (a.ncol + BUD_WARPS - 1) / BUD_WARPS;
  if (!check) {
    p3_column_conservation<Real, false><<<blocks, BUD_WARPS * 32, 0, stream>>>(a, mass, energy, rel_mass, rel_energy);
  } else {
    p3_column_conservation<Real, true><<<blocks, BUD_WARPS * 32, 0, stream>>>(a, mass, energy, rel_mass, rel_energy);
    p3_conservation_maxloc<<<1, 1024, 0, stream>>>(rel_mass, rel_energy, a.ncol, out4, (int)sizeof(Real));
  }
  return (int)cudaGetLastError();
}
template int launch_conservation<double>(const ConsArgs<double>&, bool, double*, double*, double*, double*, double*, void*);
template int launch_conservation<float>(const ConsArgs<float>&, bool, double*, double*, double*, double*, double*, void*);

template int launch_budget<double>(const MainArgs<double>&, double*, double*, int, double*, void*);
template int launch_budget<float>(const MainArgs<float>&, double*, double*, int, double*, void*);

// =================================================================================================
// The allocation-free variant of the budget (p3b200_water_energy_totals_static; no counterpart in the reference).
// The scratch of the three kernels above — col_vals[ncol][4] and the BUDGET_PARTS*4 partial sums — lives in a
// file-scope device array of a fixed size instead of the handle's cudaMalloc'ed block.  Every GPU that loads this
// library holds the array (STATIC_BUDGET_MAX_COLS*32 + BUDGET_PARTS*32 bytes = 1 MiB + 8 KiB) until the process ends.
//
// In HOST code the name g_budget_static is only a stand-in for the symbol, not a device address: the address a kernel
// can use comes from cudaGetSymbolAddress, asked once per device with that device current, and is kept per device.
// Nothing in the array has to start at zero: p3_column_budget writes all 4 values of every column < ncol before
// p3_budget_partials reads them, and every one of the BUDGET_PARTS blocks writes its partial (an empty range writes
// zeros) before p3_budget_final reads them.  The array keeps the last call's contents, and all calls on a device share
// it: two calls that overlap on the GPU (different streams, no ordering between them) race.
static __device__ __align__(256) double g_budget_static[(size_t)STATIC_BUDGET_MAX_COLS * 4 + BUDGET_PARTS * 4];

namespace {
constexpr int STATIC_MAX_DEVICES = 64;
std::mutex g_static_mutex;
double* g_static_addr[STATIC_MAX_DEVICES] = {};   // device address of g_budget_static, per device ordinal
}  // namespace

// Returns a cudaError_t as int.  The current device is the one the caller (the ABI entry point) has selected.
template <typename Real>
int launch_budget_static(const MainArgs<Real>& a, double* totals4, void* stream) {
  if (a.ncol > STATIC_BUDGET_MAX_COLS) return (int)cudaErrorInvalidValue;   // the entry point has rejected it already
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  if (dev < 0 || dev >= STATIC_MAX_DEVICES) return (int)cudaErrorInvalidDevice;
  double* base = nullptr;
  {
    std::lock_guard<std::mutex> lock(g_static_mutex);
    if (!g_static_addr[dev]) {
      void* p = nullptr;
      e = cudaGetSymbolAddress(&p, g_budget_static);
      if (e != cudaSuccess) return (int)e;
      g_static_addr[dev] = static_cast<double*>(p);
    }
    base = g_static_addr[dev];
  }
  return launch_budget<Real>(a, base, base + (size_t)STATIC_BUDGET_MAX_COLS * 4, BUDGET_PARTS, totals4, stream);
}
template int launch_budget_static<double>(const MainArgs<double>&, double*, void*);
template int launch_budget_static<float>(const MainArgs<float>&, double*, void*);

}  // namespace p3b200
