// p3_kernel.cuh — launch interface between the C ABI (p3_abi.cpp, host-only C++) and the sm_100a
// kernel translation unit (p3_kernel.cu).  Plain C++: no CUDA headers needed by includers.
#ifndef P3B200_KERNEL_CUH
#define P3B200_KERNEL_CUH

#include "p3_column.cuh"

namespace p3b200 {

constexpr int NUM_SED_CLASSES = 4;   // register-window classes of the sedimentation launches: 1, 2, 4, 8 slots per lane

struct LaunchConfig {
  int a_blocks, a_tile_cols;   // phase A: one CTA per tile of a_tile_cols columns
  unsigned a_div_magic;        // (idx * magic) >> 24 == idx / nk inside a tile
  int b_blocks[3][NUM_SED_CLASSES];   // sedimentation [cloud, rain, ice][JW = 1,2,4,8]: persistent grids (0 = class unused)
  int b_chunk;              // columns per work-queue grab (<= 32)
  int c_blocks;             // phase C: one thread per (column, level)
};

// Size the five launches for `ncol` columns of `nk` levels on the current device.  Returns a cudaError_t as int.
template <typename Real>
int plan_launch(int nk, int ncol, LaunchConfig* cfg);

// Enqueue p3_main for all columns of `a` on `stream` (a cudaStream_t).  `a.queue` must point to
// 3*NUM_SED_CLASSES zeroed device counters.  Returns a cudaError_t as int.
// `marks`: optional NUM_PHASE_MARKS cudaEvent_t: [0] before phase A, [1] after phase A, [2] cloud / [3] rain / [4] ice
// sedimentation finished, [5] all three joined (phase C may start), [6] after phase C.
// `side`: optional two helper streams + three events owned by the caller.  The three species share no data except
// precip_liq_surf (cloud and rain both add into it, atomically; phase A zeroes it): with `side` the ice and cloud
// launches run on the helper streams beside the rain launches (rain is bound by instruction issue, ice by latency),
// joined by events before phase C.  Without it the launches are serial on `stream`.
// `post`: optional p3_postamble arguments (p3_prepost.cuh); when given, phase C runs the postamble of each
// (column, level) as its tail and no separate postamble launch is needed.
constexpr int NUM_PHASE_MARKS = 7;
struct SideStreams { void* stream[2]; void* fork; void* join[2]; };   // cudaStream_t x2, cudaEvent_t x3
template <typename Real> struct PostArgs;
template <typename Real>
int launch_main(const MainArgs<Real>& a, const LaunchConfig& cfg, void* stream, void* const* marks,
                const PostArgs<Real>* post = nullptr, const SideStreams* side = nullptr);

// Water/energy budget (p3_budget.cu): per-column values into col_vals[ncol][4], fixed-order two-level sum into
// totals4 (device).  `partials` holds nparts*4 doubles of scratch.  Returns a cudaError_t as int.
constexpr int BUDGET_PARTS = 256;
template <typename Real>
int launch_budget(const MainArgs<Real>& a, double* col_vals, double* partials, int nparts, double* totals4, void* stream);

}  // namespace p3b200

// p3_prepost.cu: p3_preamble / p3_postamble launches from the C-ABI structs.  Return a cudaError_t as int.
struct P3B200PreambleFields;
struct P3B200PostambleFields;
struct P3B200CheckFields;
namespace p3b200 {
template <typename Real> int preamble_from_abi(const P3B200PreambleFields* f, int ncol, int nlev, void* stream);
// check_mask: null, or the violation mask word (p3_checks.cuh) to fold the property checks into the pass
template <typename Real> int postamble_from_abi(const P3B200PostambleFields* f, int ncol, int nlev, void* stream,
                                                unsigned* check_mask);
int checks_reset(unsigned long long* state, int real_bytes, void* stream);
template <typename Real> int checks_from_abi(const P3B200CheckFields* f, int ncol, int nlev, void* stream,
                                             unsigned long long* state);
template <typename Real> int find_loc(const void* field, int64_t stride, int ncol, int nk, double target,
                                      unsigned long long* loc_dev, void* stream);
template <typename Real> int clip_field(void* field, int64_t stride, int ncol, int nk, double lb, double ub, void* stream);
}  // namespace p3b200
#endif
