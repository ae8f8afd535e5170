/* Known-answer check for Phys::div_by_count (scream_b200/csrc/p3_physics.cuh): dt_left / n formed as
 *   y = RN(1/n); q0 = a*y; r = fma(-n, q0, a); q = fma(r, y, q0)
 * must equal the IEEE quotient a / n bit for bit for every sub-step count n the table holds, on operands in the range
 * the kernel admits (1e-250 < a < 1e250).  TEST INFRASTRUCTURE: compiled and run by tests/test_cpu_div_by_count.py.
 * usage: div_by_count_check <max n> <samples per n>; prints "checked N mismatches M", exit code 1 on a mismatch. */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

static uint64_t s = 0x9E3779B97F4A7C15ull;
static uint64_t rnd(void) { s ^= s << 13; s ^= s >> 7; s ^= s << 17; return s; }

int main(int argc, char** argv) {
  const int nmax = argc > 1 ? atoi(argv[1]) : 1023;
  const int samples = argc > 2 ? atoi(argv[2]) : 20000;
  long bad = 0, tot = 0;
  for (int n = 1; n <= nmax; ++n) {
    const double b = (double)n, y = 1.0 / b;
    for (int i = 0; i < samples; ++i) {
      const uint64_t u = rnd();
      double a;
      if (i % 3 == 0) {            /* time-step-like values, 1e-3 .. 1e5 */
        a = ldexp((double)(u >> 11) * 0x1p-53 + 0.5, (int)(rnd() % 27) - 10);
      } else if (i % 3 == 1) {     /* any exponent inside the admitted range */
        uint64_t bits = (u & 0x000FFFFFFFFFFFFFull) | ((uint64_t)(1023 - 830 + rnd() % 1660) << 52);
        memcpy(&a, &bits, 8);
      } else {                     /* multiples and near-multiples of n: exact and nearly halfway quotients */
        a = (double)(1 + (u >> 20) % 100000) * b * 0x1p-20 + (double)(rnd() % 3) * 0x1p-60;
      }
      if (!(a > 1e-250 && a < 1e250)) continue;
      const double q0 = a * y;
      const double r = fma(-b, q0, a);
      const double q = fma(r, y, q0);
      ++tot;
      if (!(q == a / b)) { if (bad < 10) printf("n=%d a=%a q=%a ref=%a\n", n, a, q, a / b); ++bad; }
    }
  }
  printf("checked %ld mismatches %ld\n", tot, bad);
  return bad != 0;
}
