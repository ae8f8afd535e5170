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

template int launch_budget<double>(const MainArgs<double>&, double*, double*, int, double*, void*);
template int launch_budget<float>(const MainArgs<float>&, double*, double*, int, double*, void*);

}  // namespace p3b200
