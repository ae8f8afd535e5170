// p3_kernel.cuh — launch interface between the C ABI (p3_abi.cpp, host-only C++) and the sm_100a
// kernel translation unit (p3_kernel.cu).  Plain C++: no CUDA headers needed by includers.
#ifndef P3B200_KERNEL_CUH
#define P3B200_KERNEL_CUH

#include "p3_column.cuh"

namespace p3b200 {

#if !defined(P3B200_A_TILE_LEVELS)
#define P3B200_A_TILE_LEVELS 4096
#endif
constexpr int A_TILE_LEVELS = P3B200_A_TILE_LEVELS;   // capacity of a tile's level list (one tile = a_tile_cols columns x nk levels)

// Bytes of phase-A scratch p3_main needs for ncol columns of nk levels in tiles of tile_cols columns
// (MainArgs::wl_idx | wl_off | lvlflags, in this order, each 256-byte aligned).
struct ScratchLayout { size_t wl_idx, wl_off, lvlflags, total; };
inline ScratchLayout scratch_layout(int ncol, int nk, int tile_cols) {
  auto up = [](size_t x) { return (x + 255) / 256 * 256; };
  const size_t ntiles = ((size_t)ncol + tile_cols - 1) / tile_cols;
  ScratchLayout L;
  L.wl_idx = 0;
  L.wl_off = up(ntiles * A_TILE_LEVELS * sizeof(unsigned short));
  L.lvlflags = L.wl_off + up(ntiles * A_OFF_STRIDE * sizeof(int));
  L.total = L.lvlflags + up((size_t)ncol * nk + 16);
  return L;
}

constexpr int NUM_SED_CLASSES = 4;   // register-window classes of the sedimentation launches: 1, 2, 4, 8 slots per lane

struct LaunchConfig {
  int a_blocks, a_tile_cols;   // phase A: one CTA per tile of a_tile_cols columns
  unsigned a_div_magic;        // (idx * magic) >> 24 == idx / nk inside a tile
  int p2_blocks;               // part2: one warp per tile, same grid for every process group's launch
  int fin_blocks;              // column finish: one thread per column
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
// `side`: optional helper streams + events owned by the caller (the six part2 group launches run side by side on
// them and the caller's stream; sedimentation uses the first two).  The three species share no data except
// precip_liq_surf (cloud and rain both add into it, atomically; phase A zeroes it): with `side` the ice and cloud
// launches run on the helper streams beside the rain launches (rain is bound by instruction issue, ice by latency),
// joined by events before phase C.  Without it the launches are serial on `stream`.
// `scratch_free`: optional cudaEvent_t recorded when the phase-A scratch (MainArgs::wl_idx/wl_off/lvlflags) is no
// longer needed by this call.
// `post`: optional p3_postamble arguments (p3_prepost.cuh); when given, phase C runs the postamble of each
// (column, level) as its tail and no separate postamble launch is needed.
// [7] after the classification pass, [8] after the part2 launches (inside phase A).
constexpr int NUM_PHASE_MARKS = 9;
constexpr int PHASE_MARK_END = 6;

constexpr int NUM_SIDE_STREAMS = 5;
struct SideStreams { void* stream[NUM_SIDE_STREAMS]; void* fork; void* join[NUM_SIDE_STREAMS]; };   // cudaStream_t, cudaEvent_t
template <typename Real> struct PostArgs;
template <typename Real>
int launch_main(const MainArgs<Real>& a, const LaunchConfig& cfg, void* stream, void* const* marks,
                const PostArgs<Real>* post = nullptr, const SideStreams* side = nullptr, void* scratch_free = nullptr);

// Water/energy budget (p3_budget.cu): per-column values into col_vals[ncol][4], fixed-order two-level sum into
// totals4 (device).  `partials` holds nparts*4 doubles of scratch.  Returns a cudaError_t as int.
constexpr int BUDGET_PARTS = 256;
template <typename Real>
int launch_budget(const MainArgs<Real>& a, double* col_vals, double* partials, int nparts, double* totals4, void* stream);
// The same three kernels with their scratch in a file-scope __device__ array of p3_budget.cu (no allocation anywhere);
// at most STATIC_BUDGET_MAX_COLS columns (= P3B200_STATIC_BUDGET_MAX_COLS of the public header).
constexpr int STATIC_BUDGET_MAX_COLS = 32768;
template <typename Real>
int launch_budget_static(const MainArgs<Real>& a, double* totals4, void* stream);

// MassAndEnergyColumnConservationCheck (p3_budget.cu).  check = false: compute_current_mass / compute_current_energy
// into mass[ncol] / energy[ncol]; check = true: relative errors per column into rel_mass / rel_energy and their MaxLoc
// into out4 = {max mass error, column, max energy error, column}.
template <typename Real>
struct ConsArgs {
  const Real *pseudo_density, *T_mid, *horiz_winds, *qv, *qc, *qi, *qr, *ps, *phis;
  const Real *vapor_flux, *water_flux, *ice_flux, *heat_flux;
  int64_t col_stride, wind_col_stride, wind_comp_stride;
  int ncol, nk;
  Real dt;
};
template <typename Real>
int launch_conservation(const ConsArgs<Real>& a, bool check, double* mass, double* energy, double* rel_mass,
                        double* rel_energy, double* out4, void* stream);

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
