// p3_kernel.cu — the sm_100a kernel of libp3b200: p3_main for a batch of columns.
//
// Execution model (replaces the Kokkos TeamPolicy of p3_main_impl.hpp:86-113, one team per column):
//   - one WARP owns one column at a time; the 32 lanes stride over the levels (lane l owns levels
//     l, l+32, ...), so every global access of a warp is one contiguous 256-byte (fp64) row segment;
//   - warps are persistent and independent: each pulls the next column index from a global atomic
//     counter, which absorbs the load imbalance of the CFL sub-stepping (columns with rain need tens
//     of sub-steps, clear-sky columns none) without sorting;
//   - the column working set (16 rows of nk Reals, ColumnSmem) lives in the warp's private slice of
//     dynamic shared memory; cross-lane traffic is warp shuffles + __syncwarp only, no __syncthreads;
//   - the lookup tables (0.65 MB fp64) are read through the read-only path and stay L2/L1 resident.
#include "p3_kernel.cuh"

#include <cuda_runtime.h>

namespace p3b200 {

struct WarpDev {
  static constexpr int size = 32;
  static constexpr unsigned FULL = 0xffffffffu;
  __device__ __forceinline__ int lane() const { return threadIdx.x & 31; }
  __device__ __forceinline__ bool any(bool b) const { return __any_sync(FULL, b) != 0; }
  __device__ __forceinline__ double max_real(double v) const {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { const double u = __shfl_xor_sync(FULL, v, o); v = (v < u) ? u : v; }
    return v;
  }
  __device__ __forceinline__ float max_real(float v) const {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { const float u = __shfl_xor_sync(FULL, v, o); v = (v < u) ? u : v; }
    return v;
  }
  __device__ __forceinline__ int min_int(int v) const { return __reduce_min_sync(FULL, v); }
  __device__ __forceinline__ int max_int(int v) const { return __reduce_max_sync(FULL, v); }
  __device__ __forceinline__ int or_int(int v) const { return (int)__reduce_or_sync(FULL, (unsigned)v); }
  __device__ __forceinline__ void sync() const { __syncwarp(); }
  // value of the previous lane (lane 0 keeps its own: callers patch it) / of lane `src`
  template <class R> __device__ __forceinline__ R shfl_up1(R v) const { return __shfl_up_sync(FULL, v, 1); }
  template <class R> __device__ __forceinline__ R shfl_from(R v, int src) const { return __shfl_sync(FULL, v, src); }
};

template <typename Real, int JM>
__global__ void __launch_bounds__(128)
p3_main_kernel(const __grid_constant__ MainArgs<Real> a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int warp = threadIdx.x >> 5;
  Real* smem = reinterpret_cast<Real*>(smem_raw) + size_t(warp) * NUM_SMEM_ROWS * a.nk;
  WarpDev w;
  const int lane = w.lane();
  for (;;) {
    unsigned int col = 0;
    if (lane == 0) col = atomicAdd(a.queue, 1u);
    col = __shfl_sync(WarpDev::FULL, col, 0);
    if (col >= (unsigned)a.ncol) break;
    Column<Real, WarpDev, JM>::run(w, a, (int64_t)col, smem);
  }
}

template <typename Real, int JM>
static int plan_for(int nk, int ncol, int sms, int max_optin, LaunchConfig* cfg) {
  const int per_warp = NUM_SMEM_ROWS * nk * (int)sizeof(Real);
  int wpb = 4;
  while (wpb > 1 && wpb * per_warp > max_optin) wpb >>= 1;
  if (wpb * per_warp > max_optin) return (int)cudaErrorInvalidValue;
  cfg->warps_per_block = wpb;
  cfg->smem_bytes = wpb * per_warp;
  cudaError_t e = cudaFuncSetAttribute(p3_main_kernel<Real, JM>, cudaFuncAttributeMaxDynamicSharedMemorySize, cfg->smem_bytes);
  if (e != cudaSuccess) return (int)e;
  int bps = 0;
  e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, p3_main_kernel<Real, JM>, wpb * 32, cfg->smem_bytes);
  if (e != cudaSuccess) return (int)e;
  if (bps < 1) return (int)cudaErrorLaunchOutOfResources;
  cfg->blocks_per_sm = bps;
  long long blocks = (long long)sms * bps;
  const long long need = ((long long)ncol + wpb - 1) / wpb;
  if (blocks > need) blocks = need;
  if (blocks < 1) blocks = 1;
  cfg->blocks = (int)blocks;
  return 0;
}

// Two instantiations per precision: JM = 4 register slots per lane (nk <= 128, the production grids)
// and JM = 8 (nk <= 256).
template <typename Real>
int plan_launch(int nk, int ncol, LaunchConfig* cfg) {
  int dev = 0, sms = 0, max_optin = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  if (e != cudaSuccess) return (int)e;
  e = cudaDeviceGetAttribute(&max_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
  if (e != cudaSuccess) return (int)e;
  if (nk > MAX_NLEV) return (int)cudaErrorInvalidValue;
  return nk <= 128 ? plan_for<Real, 4>(nk, ncol, sms, max_optin, cfg) : plan_for<Real, 8>(nk, ncol, sms, max_optin, cfg);
}

template <typename Real>
int launch_main(const MainArgs<Real>& a, const LaunchConfig& cfg, void* stream) {
  if (a.nk <= 128)
    p3_main_kernel<Real, 4><<<cfg.blocks, cfg.warps_per_block * 32, cfg.smem_bytes, (cudaStream_t)stream>>>(a);
  else
    p3_main_kernel<Real, 8><<<cfg.blocks, cfg.warps_per_block * 32, cfg.smem_bytes, (cudaStream_t)stream>>>(a);
  return (int)cudaGetLastError();
}

template int plan_launch<double>(int, int, LaunchConfig*);
template int plan_launch<float>(int, int, LaunchConfig*);
template int launch_main<double>(const MainArgs<double>&, const LaunchConfig&, void*);
template int launch_main<float>(const MainArgs<float>&, const LaunchConfig&, void*);

}  // namespace p3b200
