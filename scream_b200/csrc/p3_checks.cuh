// p3_checks.cuh — the property checks P3Microphysics registers (eamxx_p3_process_interface.cpp:226-243), fused.
//
// Reference: each FieldWithinIntervalCheck / FieldLowerBoundCheck is its own Kokkos MinMaxLoc parallel_reduce over
// a full field (field_within_interval_check.cpp:98-198) followed by the host-side comparison against the bounds
// (:202-226): 15 more passes over HBM after every P3 step.  Two pieces replace them:
//
//  * CheckFlags — rides on the postamble, which already holds all 15 values in registers: one 32-bit mask per
//    thread (bit 2i = "some value of check i is below lb", bit 2i+1 = "above ub"), OR-folded per warp and published
//    with an atomicOr only when non-zero.  No extra HBM traffic, ~one register, no atomics when everything passes.
//    A check passes iff min >= lb && max <= ub iff no value is below lb or above ub, so the mask decides
//    Pass / not-Pass exactly; NaN compares false both ways and is skipped like `v < min` / `v > max` skip it.
//  * p3_checks_kernel (p3_prepost.cu) — all 15 (min, max) pairs in ONE launch (blockIdx.y = check), for a caller
//    without the fused postamble and as the slow path when the mask is non-zero (the reference's failure message
//    needs the extreme values and where they are).  Extrema are kept as order-preserving INTEGER keys.
// Locations are found afterwards with p3_find_loc_kernel, on the failing fields only.
#pragma once
#include <cstdint>
#include <cfloat>

#if defined(__CUDACC__)
#define P3_CHK_HD __host__ __device__
#else
#define P3_CHK_HD
#endif

namespace p3b200 {

// Order of the checks — the order of the add_*_check calls in initialize_impl (:226-243).
enum CheckId { CHK_T_MID, CHK_QV, CHK_QC, CHK_QI, CHK_QR, CHK_QM, CHK_NC, CHK_NR, CHK_NI, CHK_BM,
               CHK_PRECIP_LIQ, CHK_PRECIP_ICE, CHK_RAD_QC, CHK_RAD_QI, CHK_RAD_QR, CHK_NUM };

// The registered bounds, same order; FieldLowerBoundCheck sets the upper bound to DBL_MAX
// (field_lower_bound_check.hpp:23-26).  Only qv is registered with can_repair.
constexpr double CHK_LB[CHK_NUM] = {100.0, 1e-13, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0, 0.0};
constexpr double CHK_UB[CHK_NUM] = {500.0, 0.2, 0.1, 0.1, 0.1, 0.1, 1.e11, 1.e10, 1.e10, 1.0, DBL_MAX, DBL_MAX,
                                    1.0e2, 5.0e3, 5.0e3};
constexpr bool CHK_CAN_REPAIR[CHK_NUM] = {false, true, false, false, false, false, false, false, false, false,
                                          false, false, false, false, false};
constexpr bool CHK_COLUMN_FIELD[CHK_NUM] = {false, false, false, false, false, false, false, false, false, false,
                                            true, true, false, false, false};
// device words behind the handle: 2*CHK_NUM min/max keys, 2 location slots, 1 violation mask
constexpr int CHK_STATE_LOC = 2 * CHK_NUM, CHK_STATE_MASK = 2 * CHK_NUM + 2, CHK_STATE_WORDS = 2 * CHK_NUM + 3;

// monotone map double -> u64 (so that integer atomicMin/Max order like the floating-point values)
P3_CHK_HD inline unsigned long long chk_encode(double v) {
  unsigned long long u;
#if defined(__CUDA_ARCH__)
  u = (unsigned long long)__double_as_longlong(v);
#else
  static_assert(sizeof(double) == sizeof(u), "");
  __builtin_memcpy(&u, &v, 8);
#endif
  return (u >> 63) ? ~u : (u | 0x8000000000000000ull);
}
P3_CHK_HD inline double chk_decode(unsigned long long u) {
  u = (u >> 63) ? (u & 0x7fffffffffffffffull) : ~u;
  double v;
#if defined(__CUDA_ARCH__)
  v = __longlong_as_double((long long)u);
#else
  __builtin_memcpy(&v, &u, 8);
#endif
  return v;
}

#if defined(__CUDACC__)
// Order-preserving integer keys of the field's own type: the running extrema are kept as INTEGERS, so that the
// reduction never touches the FP64 pipe and NaN is excluded by an integer range test.
template <typename Real> struct CheckKey;
template <> struct CheckKey<double> {
  using K = unsigned long long;
  static constexpr K SIGN = 0x8000000000000000ull, INF = 0x7ff0000000000000ull;
  static __device__ __forceinline__ K bits(double v) { return (K)__double_as_longlong(v); }
  static __device__ __forceinline__ double value(K u) { return __longlong_as_double((long long)u); }
  static __device__ __forceinline__ K big() { return bits(DBL_MAX); }
};
template <> struct CheckKey<float> {
  using K = unsigned int;
  static constexpr K SIGN = 0x80000000u, INF = 0x7f800000u;
  static __device__ __forceinline__ K bits(float v) { return __float_as_uint(v); }
  static __device__ __forceinline__ float value(K u) { return __uint_as_float(u); }
  static __device__ __forceinline__ K big() { return bits(FLT_MAX); }
};

// The violation mask of one thread (fused into the postamble).
template <typename Real>
struct CheckFlags {
  unsigned mask;
  __device__ __forceinline__ void init() { mask = 0u; }
  template <int ID>
  __device__ __forceinline__ void see(Real v) {
    const double d = (double)v;                 // the reference compares the Real extreme with double bounds
    constexpr double lb = CHK_LB[ID], ub = CHK_UB[ID];
    if (d < lb) mask |= 1u << (2 * ID);
    if (d > ub) mask |= 2u << (2 * ID);
  }
  // called by ALL threads of the (converged) block
  __device__ __forceinline__ void publish(unsigned* word) {
    const unsigned m = __reduce_or_sync(0xffffffffu, mask);
    if ((threadIdx.x & 31) == 0 && m) atomicOr(word, m);
  }
};

// Running (min, max) of ONE field as integer keys.  Kokkos' reduction identities: min starts at the largest finite
// Real, max at its negative (Kokkos::reduction_identity<Real>::min()/max()).
template <typename Real>
struct CheckMinMax {
  using CK = CheckKey<Real>;
  using K = typename CK::K;
  K mn, mx;
  static __device__ __forceinline__ K encode(K u) { return (u & CK::SIGN) ? ~u : (u | CK::SIGN); }
  static __device__ __forceinline__ K decode(K k) { return (k & CK::SIGN) ? (k & ~CK::SIGN) : ~k; }
  static __device__ __forceinline__ Real skipped() { return CK::value(~K(0)); }      // a NaN: never wins
  __device__ __forceinline__ void init() { mn = encode(CK::big()); mx = encode(CK::big() | CK::SIGN); }
  __device__ __forceinline__ void see(Real v) {
    const K u = CK::bits(v);
    const bool is_nan = (u & ~CK::SIGN) > CK::INF;      // NaN never wins `v < min` / `v > max`
    const K k = encode(u);
    const K lo = is_nan ? ~K(0) : k, hi = is_nan ? K(0) : k;
    mn = lo < mn ? lo : mn;
    mx = hi > mx ? hi : mx;
  }
  // Fold the block and publish 64-bit keys of the values as doubles (chk_encode): state[0] min, state[1] max.
  template <int THREADS>
  __device__ __forceinline__ void publish(unsigned long long* state) {
    __shared__ K s_red[THREADS / 32][2];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    K a = mn, b = mx;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const K a2 = __shfl_xor_sync(0xffffffffu, a, o), b2 = __shfl_xor_sync(0xffffffffu, b, o);
      a = a2 < a ? a2 : a;
      b = b2 > b ? b2 : b;
    }
    if (lane == 0) { s_red[warp][0] = a; s_red[warp][1] = b; }
    __syncthreads();
    if (threadIdx.x == 0) {
      for (int w = 1; w < THREADS / 32; ++w) {
        a = s_red[w][0] < a ? s_red[w][0] : a;
        b = s_red[w][1] > b ? s_red[w][1] : b;
      }
      atomicMin(state, chk_encode((double)CK::value(decode(a))));
      atomicMax(state + 1, chk_encode((double)CK::value(decode(b))));
    }
  }
};
#endif

}  // namespace p3b200
