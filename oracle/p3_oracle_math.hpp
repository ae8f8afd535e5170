// ORACLE — TEST INFRASTRUCTURE ONLY.
// Transcendental functions of the restatement, routed through one place so that the tests can ask
// "how much of a result is decided by the last bit of libm?".  The reference itself states that its
// CPU and GPU builds are not bit-for-bit because std:: and CUDA math differ
// (physics/share/physics_share.cpp:17; baselines are per platform), and P3 is a cascade
// of thresholds (qsmall gates, int(Co_max+1) sub-step counts, table-cell truncation), so a 1-ulp
// difference can be amplified without bound in some columns.  With every scale == 1.0 (the default)
// each wrapper returns the libm value bit for bit (x * 1.0 is exact); the conditioning probe of
// tests/parity.py sets them to 1 +- 2^-52 to find the columns whose outcome depends on that last bit.
#ifndef P3O_MATH_HPP
#define P3O_MATH_HPP
//
// Two math back ends, chosen at run time (p3o_set_math_mode): 0 = the host libm (std::, what the reference's CPU
// build calls), 1 = the deterministic transcendentals of scream_b200/csrc/p3_math.hpp, the header the sm_100a
// kernels include too — the reference's own device for bit-for-bit comparisons between two builds
// (physics/share/physics_share.cpp:17-150), here between the CPU oracle and the GPU.
#include <cmath>

#include "../scream_b200/csrc/p3_math.hpp"

namespace p3o {
namespace pm {
enum Fn { POW = 0, EXP, LOG, LOG10, CBRT, TGAMMA, TANH, EXPM1, NUM_FN };
extern double g_scale[NUM_FN];   // defined in p3_oracle.cpp; written only between runs
extern int g_shared;             // 0: libm, 1: p3math (shared with the kernels)

template <class R> inline R pow(R x, R y)  { return R((g_shared ? p3math::pow(x, y) : std::pow(x, y)) * R(g_scale[POW])); }
template <class R> inline R exp(R x)       { return R((g_shared ? p3math::exp(x) : std::exp(x)) * R(g_scale[EXP])); }
template <class R> inline R log(R x)       { return R((g_shared ? p3math::log(x) : std::log(x)) * R(g_scale[LOG])); }
template <class R> inline R log10(R x)     { return R((g_shared ? p3math::log10(x) : std::log10(x)) * R(g_scale[LOG10])); }
template <class R> inline R cbrt(R x)      { return R((g_shared ? p3math::cbrt(x) : std::cbrt(x)) * R(g_scale[CBRT])); }
template <class R> inline R tgamma(R x)    { return R((g_shared ? p3math::tgamma(x) : std::tgamma(x)) * R(g_scale[TGAMMA])); }
template <class R> inline R tanh(R x)      { return R((g_shared ? p3math::tanh(x) : std::tanh(x)) * R(g_scale[TANH])); }
template <class R> inline R expm1(R x)     { return R((g_shared ? p3math::expm1(x) : std::expm1(x)) * R(g_scale[EXPM1])); }
}  // namespace pm
}  // namespace p3o
#endif
