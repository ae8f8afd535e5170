// p3_math.hpp — deterministic transcendentals shared VERBATIM by the sm_100a kernels and the CPU oracle.
//
// Why: P3 is a cascade of thresholds (qsmall gates, int(Co_max+1) sub-step counts, table-cell truncation) and a
// 1-ulp difference between glibc's libm and CUDA's libdevice flips such a threshold in ~13 % of random columns
// (DESIGN.md §4).  The reference meets the same problem between its Fortran and C++ builds and solves it by routing
// both through ONE implementation of pow/exp/log/... (components/eamxx/src/physics/share/physics_share.cpp:17-150,
// scream_pow, scream_gamma, scream_cbrt, ...).  This header is that idea for the CPU<->GPU pair: every function below
// uses only IEEE-754 binary64 + - * / fma, integer operations on the bit pattern and comparisons — no libm, no
// libdevice, no SFU approximations — so a host compiler (g++ -ffp-contract=off, hardware or software fma) and nvcc
// (-fmad=false; fma() is always fused) produce bit-identical results.  Accuracy is in the class of a good libm
// (measured against mpmath in tests/test_cpu_shared_math.py: exp/log/log10/cbrt/expm1/tanh < 1 ulp, pow < 1 ulp for
// |y log x| < 700, tgamma <= 5 ulp on (0, 171]).
//
// The float overloads evaluate in binary64 and round once: deterministic, and closer to the exact value than a
// native binary32 routine.
#ifndef P3B200_MATH_HPP
#define P3B200_MATH_HPP

#include <cmath>
#include <cstdint>
#include <cstring>

#if defined(__CUDACC__)
#define P3M_HD __host__ __device__ __forceinline__
#define P3M_HD_NOINLINE inline __host__ __device__ __noinline__
#else
#define P3M_HD inline
#define P3M_HD_NOINLINE inline
#endif

namespace p3math {

P3M_HD int64_t d2i(double x) {
#if defined(__CUDA_ARCH__)
  return __double_as_longlong(x);
#else
  int64_t i; std::memcpy(&i, &x, 8); return i;
#endif
}
P3M_HD double i2d(int64_t i) {
#if defined(__CUDA_ARCH__)
  return __longlong_as_double(i);
#else
  double x; std::memcpy(&x, &i, 8); return x;
#endif
}
P3M_HD double fma_(double a, double b, double c) {
#if defined(__CUDA_ARCH__)
  return __fma_rn(a, b, c);
#else
  return std::fma(a, b, c);
#endif
}
P3M_HD double inf_() { return i2d(0x7ff0000000000000LL); }
P3M_HD double nan_() { return i2d(0x7ff8000000000000LL); }
P3M_HD double abs_(double x) { return i2d(d2i(x) & 0x7fffffffffffffffLL); }
P3M_HD double copysign_(double mag, double sgn) {
  return i2d((d2i(mag) & 0x7fffffffffffffffLL) | (d2i(sgn) & (int64_t)0x8000000000000000ULL));
}
// p * 2^k for |k| <= 2100 with a single rounding (the first factor keeps the product in the normal range)
P3M_HD double scale2(double p, int k) {
  const int k1 = k / 2, k2 = k - k1;
  return p * i2d((int64_t)(k1 + 1023) << 52) * i2d((int64_t)(k2 + 1023) << 52);
}

struct dd { double hi, lo; };

constexpr double LN2_HI32 = 0x1.62e42fee00000p-1;   // ln 2, 32 significant bits: k * LN2_HI32 is exact for |k| < 2^21
constexpr double LN2_LO32 = 0x1.a39ef35793c76p-33;
constexpr double LN2_HI42 = 0x1.62e42fefa3800p-1;   // 42 significant bits: e * LN2_HI42 is exact for |e| < 2^11
constexpr double LN2_LO42 = 0x1.ef35793c76730p-45;
constexpr double INV_LN2  = 0x1.71547652b82fep+0;
constexpr double RINT_SHIFT = 0x1.8p52;

// exp(hi + lo) for |lo| << |hi| (lo: the low word of a double-double exponent).  k = rint(hi / ln 2), Cody-Waite
// reduction with an exact first step, Taylor polynomial of degree 13 on |r| <= 0.3466 (truncation 4e-18) in Estrin form.
P3M_HD double exp_dd(double hi, double lo) {
  if (!(hi <= 709.79)) return (hi != hi) ? hi : inf_();     // NaN, +inf, overflow
  if (hi < -745.14) return 0.0;
  const double kd = (hi * INV_LN2 + RINT_SHIFT) - RINT_SHIFT;
  double r = fma_(-kd, LN2_HI32, hi);
  r = fma_(-kd, LN2_LO32, r);
  r = r + lo;
  const double r2 = r * r, r4 = r2 * r2, r8 = r4 * r4;
  // q(r) = sum_{i=2..13} r^(i-2) / i!
  const double p01 = fma_(r, 0x1.5555555555555p-3, 0x1.0000000000000p-1);
  const double p23 = fma_(r, 0x1.1111111111111p-7, 0x1.5555555555555p-5);
  const double p45 = fma_(r, 0x1.a01a01a01a01ap-13, 0x1.6c16c16c16c17p-10);
  const double p67 = fma_(r, 0x1.71de3a556c734p-19, 0x1.a01a01a01a01ap-16);
  const double p89 = fma_(r, 0x1.ae64567f544e4p-26, 0x1.27e4fb7789f5cp-22);
  const double pab = fma_(r, 0x1.6124613a86d09p-33, 0x1.1eed8eff8d898p-29);
  const double q0 = fma_(r2, p23, p01), q1 = fma_(r2, p67, p45), q2 = fma_(r2, pab, p89);
  const double q = fma_(r8, q2, fma_(r4, q1, q0));
  const double p = 1.0 + fma_(r2, q, r);
  return scale2(p, (int)kd);
}

// log(x) as a double-double (relative error ~2^-66) for finite x > 0.  x = 2^e m, m in [sqrt(1/2), sqrt(2));
// s = (m-1)/(m+1) as a double-double; log m = 2s + s*(2/3 z + 2 z^2 g(z)), z = s^2, g a degree-7 near-minimax fit
// (mpmath.chebyfit) of (atanh(s)/s - 1 - z/3)/z^2 on [0, 0.02944].
P3M_HD dd log_dd(double x) {
  int64_t ix = d2i(x);
  int e = 0;
  if (ix < 0x0010000000000000LL) { x = x * 0x1p54; ix = d2i(x); e = -54; }   // subnormal
  const int64_t k = (ix - 0x3fe6a09e667f3bcdLL) >> 52;
  e += (int)k;
  const double m = i2d(ix - (k << 52));
  const double f = m - 1.0;                                  // exact
  const double dh = m + 1.0;                                 // TwoSum(m, 1)
  const double bb = dh - m;
  const double dl = (m - (dh - bb)) + (1.0 - bb);
  const double rcp = 1.0 / dh;
  const double sh = f * rcp;
  const double rem = fma_(-sh, dh, f);
  const double sl = (rem - sh * dl) * rcp;
  const double zh = sh * sh;
  const double zl = fma_(sh, sh, -zh) + 2.0 * sh * sl;
  double g = 0x1.e02e6fff27908p-5;
  g = fma_(g, zh, 0x1.df7cc8e5fa539p-5);
  g = fma_(g, zh, 0x1.1118c7e4b0917p-4);
  g = fma_(g, zh, 0x1.3b1395d2190c1p-4);
  g = fma_(g, zh, 0x1.745d177a96056p-4);
  g = fma_(g, zh, 0x1.c71c71c6eaa10p-4);
  g = fma_(g, zh, 0x1.2492492492521p-3);
  g = fma_(g, zh, 0x1.999999999999ap-3);
  constexpr double C3H = 0x1.5555555555555p-1, C3L = 0x1.5555555555555p-55;   // 2/3
  const double qh = zh * C3H;
  const double ql = fma_(zh, C3H, -qh) + zh * C3L + zl * C3H + (zh * zh) * (2.0 * g);
  const double ph = sh * qh;
  const double pl = fma_(sh, qh, -ph) + (sh * ql + sl * qh);
  const double a = 2.0 * sh;
  const double hi = a + ph;
  double lo = (a - hi) + ph;
  lo = lo + (pl + 2.0 * sl);
  const double ed = (double)e;
  const double eh = ed * LN2_HI42;                           // exact
  const double H = eh + hi;
  double L = (eh - H) + hi;
  L = L + (lo + ed * LN2_LO42);
  dd r;
  r.hi = H + L;
  r.lo = (H - r.hi) + L;
  return r;
}

P3M_HD double exp(double x) { return exp_dd(x, 0.0); }

P3M_HD double log(double x) {
  if (!(x > 0.0)) return (x == 0.0) ? -inf_() : nan_();      // 0, negative, NaN
  if (x == inf_()) return x;
  return log_dd(x).hi;
}

P3M_HD double log10(double x) {
  if (!(x > 0.0)) return (x == 0.0) ? -inf_() : nan_();
  if (x == inf_()) return x;
  const dd l = log_dd(x);
  constexpr double CH = 0x1.bcb7b1526e50ep-2, CL = 0x1.95355baaafad3p-57;    // 1 / ln 10
  const double h = l.hi * CH;
  const double t = fma_(l.hi, CH, -h) + (l.hi * CL + l.lo * CH);
  return h + t;
}

// pow(x, y) = exp(y * log x) with the logarithm and the product carried as double-doubles.
P3M_HD double pow(double x, double y) {
  if (y == 0.0 || x == 1.0) return 1.0;
  if (x != x || y != y) return nan_();
  double sign = 1.0;
  if (!(x > 0.0)) {
    // y integral?  (|y| >= 2^53 is an even integer)
    const double ay = abs_(y);
    const double yr = (ay < 0x1p53) ? ((ay + RINT_SHIFT) - RINT_SHIFT) : ay;
    const bool y_int = (yr == ay);
    const bool y_odd = y_int && ay < 0x1p53 && ((int64_t)ay & 1);
    if (x == 0.0) {
      const double z = (y < 0.0) ? inf_() : 0.0;
      return y_odd ? copysign_(z, x) : z;
    }
    if (!y_int) return nan_();
    if (y_odd) sign = -1.0;
    x = -x;
  }
  const double ay = abs_(y);
  if (x == inf_()) return sign * ((y < 0.0) ? 0.0 : inf_());
  if (ay == inf_()) return ((x > 1.0) == (y > 0.0)) ? inf_() : 0.0;
  const dd l = log_dd(x);
  const double ph = y * l.hi;
  const double pl = fma_(y, l.hi, -ph) + y * l.lo;
  return sign * exp_dd(ph, pl);
}

// expm1(x): k = rint(x / ln 2), expm1(r) = r + r^2 q(r) (Taylor, degree 14), result 2^k expm1(r) + (2^k - 1).
P3M_HD double expm1(double x) {
  if (x != x) return x;
  if (x > 709.79) return inf_();
  if (x < -38.0) return -1.0;
  const double kd = (x * INV_LN2 + RINT_SHIFT) - RINT_SHIFT;
  double r = fma_(-kd, LN2_HI32, x);
  r = fma_(-kd, LN2_LO32, r);
  const double r2 = r * r, r4 = r2 * r2, r8 = r4 * r4;
  const double p01 = fma_(r, 0x1.5555555555555p-3, 0x1.0000000000000p-1);
  const double p23 = fma_(r, 0x1.1111111111111p-7, 0x1.5555555555555p-5);
  const double p45 = fma_(r, 0x1.a01a01a01a01ap-13, 0x1.6c16c16c16c17p-10);
  const double p67 = fma_(r, 0x1.71de3a556c734p-19, 0x1.a01a01a01a01ap-16);
  const double p89 = fma_(r, 0x1.ae64567f544e4p-26, 0x1.27e4fb7789f5cp-22);
  const double pab = fma_(r, 0x1.6124613a86d09p-33, 0x1.1eed8eff8d898p-29);
  const double pcd = 0x1.93974a8c07c9dp-37;
  const double q0 = fma_(r2, p23, p01), q1 = fma_(r2, p67, p45), q2 = fma_(r2, pab, p89);
  const double q = fma_(r8, fma_(r4, pcd, q2), fma_(r4, q1, q0));
  const double em = fma_(r2, q, r);
  const int k = (int)kd;
  if (k == 0) return em;
  if (k > 56) return scale2(1.0 + em, k);                    // the -1 is below half an ulp
  const double tk = i2d((int64_t)(k + 1023) << 52);          // 2^k, -55 <= k <= 56
  return fma_(tk, em, tk - 1.0);
}

P3M_HD double tanh(double x) {
  if (x != x) return x;
  const double a = abs_(x);
  double z;
  if (a >= 22.0) z = 1.0;
  else if (a < 0x1p-28) z = a;
  else {
    const double t = expm1(2.0 * a);
    z = (a < 1.0) ? t / (t + 2.0) : 1.0 - 2.0 / (t + 2.0);
  }
  return copysign_(z, x);
}

// cbrt(x): x = 2^(3q+j) m; degree-6 near-minimax fit of m^(1/3) on [1,2) (2^-22) times 2^(j/3), then one Halley step
// on y = 2^j m in [1,8) with the residual y - t^3 formed exactly by fma: t <- t + t e / (3y - 2e), error < 0.51 ulp.
P3M_HD double cbrt(double x) {
  int64_t ix = d2i(x);
  const int64_t sgn = ix & (int64_t)0x8000000000000000ULL;
  ix ^= sgn;
  if (ix >= 0x7ff0000000000000LL || ix == 0) return x;       // inf, NaN, +-0
  int e = 0;
  if (ix < 0x0010000000000000LL) { ix = d2i(i2d(ix) * 0x1p54); e = -54; }
  e += (int)(ix >> 52) - 1023;
  const double m = i2d((ix & 0x000fffffffffffffLL) | 0x3ff0000000000000LL);
  const int q = (e + 3072) / 3 - 1024;                       // floor(e / 3)
  const int j = e - 3 * q;
  double t = -0x1.6254d5208eb99p-9;
  t = fma_(t, m, 0x1.e346250683378p-6);
  t = fma_(t, m, -0x1.1d2e558db1d25p-3);
  t = fma_(t, m, 0x1.8026ac948a766p-2);
  t = fma_(t, m, -0x1.4f65258278179p-1);
  t = fma_(t, m, 0x1.e33114033c8b7p-1);
  t = fma_(t, m, 0x1.cb68f94b94beap-2);
  t = t * ((j == 0) ? 1.0 : (j == 1) ? 0x1.428a2f98d728bp+0 : 0x1.965fea53d6e3dp+0);
  const double y = m * ((j == 0) ? 1.0 : (j == 1) ? 2.0 : 4.0);
  const double s = t * t;
  const double sl = fma_(t, t, -s);
  const double er = fma_(-s, t, y) - sl * t;                 // y - t^3
  const double den = fma_(-2.0, er, 3.0 * y);
  t = fma_(t, er / den, t);
  return i2d(d2i(scale2(t, q)) | sgn);
}

// tgamma(x).  x is shifted up to y >= 10 (Gamma(x) = Gamma(y) / (x (x+1) ... (y-1))), then
// Gamma(y) = exp((y - 1/2) log y - y + log(2 pi)/2 + S(1/y)) with the exponent carried as a double-double and the
// Stirling series S through y^-15 (next term < 2e-18).  The rounding error of y = x + n is put back through
// digamma(y) ~ log y - 1/(2y).
P3M_HD_NOINLINE double tgamma(double x) {
  if (x != x) return x;
  if (x == inf_()) return x;
  if (x > 171.7) return inf_();
  if (x <= 0.0) {
    if (x == 0.0) return copysign_(inf_(), x);
    const double xr = (x > -0x1p52) ? ((x - RINT_SHIFT) + RINT_SHIFT) : x;
    if (xr == x) return nan_();                              // pole
    if (x < -190.0) return 0.0;
  }
  double prod = 1.0, y = x;
  int n = 0;
  while (y < 10.0) { prod = prod * y; n += 1; y = x + (double)n; }
  // y = x + n rounded; yl = exact error of that sum (TwoSum)
  const double nd = (double)n;
  const double bb = y - x;
  const double yl = (x - (y - bb)) + (nd - bb);
  const dd l = log_dd(y);
  const double a = y - 0.5;                                  // exact for y >= 10 (y has at most 49 fraction bits)
  const double ph = a * l.hi;
  double pl = fma_(a, l.hi, -ph) + a * l.lo;
  const double w = 1.0 / y, w2 = w * w;
  double s = 0x1.a41a41a41a41ap-8;                           // B_2n / (2n (2n-1)), n = 7 .. 1
  s = fma_(s, w2, -0x1.f6ab0d9993c7dp-10);
  s = fma_(s, w2, 0x1.b951e2b18ff23p-11);
  s = fma_(s, w2, -0x1.3813813813814p-11);
  s = fma_(s, w2, 0x1.a01a01a01a01ap-11);
  s = fma_(s, w2, -0x1.6c16c16c16c17p-9);
  s = fma_(s, w2, 0x1.5555555555555p-4);
  s = s * w;
  constexpr double HL2PI_H = 0x1.d67f1c864beb5p-1, HL2PI_L = -0x1.65b5a1b7ff5dfp-55;
  // E = (ph - y) + HL2PI_H  +  [pl + s + HL2PI_L + yl * digamma(y)]
  const double t1 = ph - y;                                  // TwoSum(ph, -y)
  const double b1 = t1 - ph;
  const double e1 = (ph - (t1 - b1)) + (-y - b1);
  const double t2 = t1 + HL2PI_H;
  const double b2 = t2 - t1;
  const double e2 = (t1 - (t2 - b2)) + (HL2PI_H - b2);
  pl = pl + e1 + e2 + s + HL2PI_L + yl * (l.hi - 0.5 * w);
  const double gy = exp_dd(t2, pl);
  return gy / prod;
}

// ---- binary32 face: evaluate in binary64, round once ----
P3M_HD float exp(float x) { return (float)exp((double)x); }
P3M_HD float log(float x) { return (float)log((double)x); }
P3M_HD float log10(float x) { return (float)log10((double)x); }
P3M_HD float pow(float x, float y) { return (float)pow((double)x, (double)y); }
P3M_HD float expm1(float x) { return (float)expm1((double)x); }
P3M_HD float tanh(float x) { return (float)tanh((double)x); }
P3M_HD float cbrt(float x) { return (float)cbrt((double)x); }
P3M_HD float tgamma(float x) { return (float)tgamma((double)x); }

}  // namespace p3math
#endif
