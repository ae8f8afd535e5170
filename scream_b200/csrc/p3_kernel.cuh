// p3_kernel.cuh — launch interface between the C ABI (p3_abi.cpp, host-only C++) and the sm_100a
// kernel translation unit (p3_kernel.cu).  Plain C++: no CUDA headers needed by includers.
#ifndef P3B200_KERNEL_CUH
#define P3B200_KERNEL_CUH

#include "p3_column.cuh"

namespace p3b200 {

struct LaunchConfig {
  int warps_per_block;   // independent warps per CTA (no block-level sync is ever used)
  int blocks;            // persistent grid: SM count x resident CTAs per SM
  int smem_bytes;        // dynamic shared memory per CTA
  int blocks_per_sm;     // occupancy the config was sized for
};

// Size a persistent grid for `nk` levels on the current device.  Returns a cudaError_t as int.
template <typename Real>
int plan_launch(int nk, int ncol, LaunchConfig* cfg);

// Enqueue p3_main for all columns of `a` on `stream` (a cudaStream_t).  `a.queue` must point to a
// zeroed device counter.  Returns a cudaError_t as int.
template <typename Real>
int launch_main(const MainArgs<Real>& a, const LaunchConfig& cfg, void* stream);

}  // namespace p3b200
#endif
