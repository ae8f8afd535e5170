// p3_ref_mathhook.cpp — TEST INFRASTRUCTURE ONLY.
// Linked (with -Wl,-Bsymbolic-functions) into the parity variants of oracle/_ref: defines the C math symbols the
// reference's code calls (pow, exp, log, log10, cbrt, tgamma, tanh, expm1 and their float twins) so that they can be
// switched, at run time, between the host libm (default: the reference's CPU behaviour) and the deterministic
// transcendentals of scream_b200/csrc/p3_math.hpp — the header the sm_100a kernels and the oracle share.  With the
// switch on, the reference's OWN unmodified p3_main runs on exactly the arithmetic of the GPU parity build, which
// closes the chain  reference code == oracle == GPU kernels  bit for bit on every column.
// This is the reference's own device (physics/share/physics_share.cpp:17-150: one math implementation under both
// builds being compared), applied from the outside: no reference source is modified.
#include <dlfcn.h>

#include <cstdio>
#include <cstdlib>

#include "../../scream_b200/csrc/p3_math.hpp"

namespace {
int g_shared = 0;
struct Libm {
  double (*pow)(double, double); double (*exp)(double); double (*log)(double); double (*log10)(double);
  double (*cbrt)(double); double (*tgamma)(double); double (*tanh)(double); double (*expm1)(double);
  float (*powf)(float, float); float (*expf)(float); float (*logf)(float); float (*log10f)(float);
  float (*cbrtf)(float); float (*tgammaf)(float); float (*tanhf)(float); float (*expm1f)(float);
  Libm() {
    void* h = dlopen("libm.so.6", RTLD_NOW | RTLD_LOCAL);
    if (!h) { std::fprintf(stderr, "p3ref mathhook: cannot open libm.so.6\n"); std::abort(); }
#define P3REF_SYM(n) n = reinterpret_cast<decltype(n)>(dlsym(h, #n)); if (!n) { std::fprintf(stderr, "p3ref mathhook: no %s\n", #n); std::abort(); }
    P3REF_SYM(pow) P3REF_SYM(exp) P3REF_SYM(log) P3REF_SYM(log10) P3REF_SYM(cbrt) P3REF_SYM(tgamma) P3REF_SYM(tanh) P3REF_SYM(expm1)
    P3REF_SYM(powf) P3REF_SYM(expf) P3REF_SYM(logf) P3REF_SYM(log10f) P3REF_SYM(cbrtf) P3REF_SYM(tgammaf) P3REF_SYM(tanhf) P3REF_SYM(expm1f)
#undef P3REF_SYM
  }
};
const Libm& libm() { static Libm l; return l; }
}  // namespace

extern "C" {
void p3ref_set_math_mode(int shared) { g_shared = shared ? 1 : 0; (void)libm(); }
int p3ref_get_math_mode() { return g_shared; }
int p3ref_has_math_hook() { return 1; }

double pow(double x, double y) { return g_shared ? p3math::pow(x, y) : libm().pow(x, y); }
double exp(double x) { return g_shared ? p3math::exp(x) : libm().exp(x); }
double log(double x) { return g_shared ? p3math::log(x) : libm().log(x); }
double log10(double x) { return g_shared ? p3math::log10(x) : libm().log10(x); }
double cbrt(double x) { return g_shared ? p3math::cbrt(x) : libm().cbrt(x); }
double tgamma(double x) { return g_shared ? p3math::tgamma(x) : libm().tgamma(x); }
double tanh(double x) { return g_shared ? p3math::tanh(x) : libm().tanh(x); }
double expm1(double x) { return g_shared ? p3math::expm1(x) : libm().expm1(x); }
float powf(float x, float y) { return g_shared ? p3math::pow(x, y) : libm().powf(x, y); }
float expf(float x) { return g_shared ? p3math::exp(x) : libm().expf(x); }
float logf(float x) { return g_shared ? p3math::log(x) : libm().logf(x); }
float log10f(float x) { return g_shared ? p3math::log10(x) : libm().log10f(x); }
float cbrtf(float x) { return g_shared ? p3math::cbrt(x) : libm().cbrtf(x); }
float tgammaf(float x) { return g_shared ? p3math::tgamma(x) : libm().tgammaf(x); }
float tanhf(float x) { return g_shared ? p3math::tanh(x) : libm().tanhf(x); }
float expm1f(float x) { return g_shared ? p3math::expm1(x) : libm().expm1f(x); }
}
