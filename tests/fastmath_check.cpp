// Host accuracy check of slam-duckietown_b200/csrc/ekf_fastmath.cuh (compiled by tests/test_fastmath.py with g++;
// the MUFU seeds are replaced by float-precision stand-ins on the host, everything else is the device code).
// Prints the worst error in ulp of each function against long double libm, then special values of atan2.
#include <cstdio>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <random>
#include "ekf_fastmath.cuh"
static double ulp_err(double got, long double want) {
  if (want == 0) return got == 0 ? 0 : 1e9;
  int e; frexpl(fabsl(want), &e);
  long double ulp = ldexpl(1.0L, e - 53);
  return (double)(fabsl((long double)got - want) / ulp);
}
int main() {
  std::mt19937_64 g(1);
  std::uniform_real_distribution<double> U(-1, 1), E(-8, 8);
  double worst_a = 0, worst_d = 0, worst_r = 0, worst_s = 0, worst_is = 0; double wa_y=0, wa_x=0;
  for (long i = 0; i < 2000000; ++i) {
    double x = U(g) * pow(10, E(g) * (i & 1)), y = U(g) * pow(10, E(g) * ((i >> 1) & 1));
    if (i % 1000 == 0) y = x * 0.41421356237309503 * (1 + 1e-15 * (i % 7 - 3));
    if (i % 1000 == 1) y = x;
    if (i % 1000 == 2) y = 1e-300 * x;
    double a = ekf::fm_atan2(y, x);
    double ea = ulp_err(a, atan2l((long double)y, (long double)x));
    if (ea > worst_a) { worst_a = ea; wa_y = y; wa_x = x; }
    double d = fabs(x) + 1e-30;
    worst_d = fmax(worst_d, ulp_err(ekf::fm_div(y, d), (long double)y / d));
    worst_r = fmax(worst_r, ulp_err(ekf::fm_rcp(d), 1.0L / d));
    double r, ir; ekf::fm_sqrt_rsqrt(d, &r, &ir);
    worst_s = fmax(worst_s, ulp_err(r, sqrtl((long double)d)));
    worst_is = fmax(worst_is, ulp_err(ir, 1.0L / sqrtl((long double)d)));
  }
  double worst_sin = 0, worst_cos = 0;
  std::uniform_real_distribution<double> A(-1, 1);
  for (long i = 0; i < 2000000; ++i) {
    double x = A(g) * ((i & 3) == 0 ? 1.0e5 : (i & 3) == 1 ? 100.0 : 3.2);
    if (i % 1000 == 0) x = (double)(i / 1000 % 64) * 0.78539816339744828 * (1 + 1e-15 * (i % 5 - 2));
    double s, c; ekf::fm_sincos(x, &s, &c);
    worst_sin = fmax(worst_sin, fmin(ulp_err(s, sinl((long double)x)), fabs(s - (double)sinl((long double)x)) / 1.1e-16));
    worst_cos = fmax(worst_cos, fmin(ulp_err(c, cosl((long double)x)), fabs(c - (double)cosl((long double)x)) / 1.1e-16));
  }
  printf("sin %.3f cos %.3f ulp\n", worst_sin, worst_cos);
  printf("atan2 %.3f (y=%a x=%a) div %.3f rcp %.3f sqrt %.3f rsqrt %.3f ulp\n", worst_a, wa_y, wa_x, worst_d, worst_r, worst_s, worst_is);
  // special values
  double sp[][2] = {{0.0, 1.0}, {0.0, -1.0}, {-0.0, -1.0}, {1.0, 0.0}, {-1.0, 0.0}, {0.0, 0.0}, {0.0, -0.0}, {-0.0, -0.0}, {1.0, 1.0}, {-1.0, -1.0}, {1.0, -1.0}};
  for (auto& v : sp) printf("atan2(%g,%g) = %.17g libm %.17g\n", v[0], v[1], ekf::fm_atan2(v[0], v[1]), atan2(v[0], v[1]));
}
