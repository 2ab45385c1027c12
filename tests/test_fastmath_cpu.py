"""Accuracy of mcmc_b200/csrc/fastmath.cuh (the Box-Muller building blocks), compiled for the HOST
with g++ from the very source the kernels include, against glibc / long-double references."""
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

SRC = r'''
#include <cstdio>
#include <cmath>
#include <cstdint>
#include <random>
#include "%s/mcmc_b200/csrc/fastmath.cuh"
using namespace mcmc;
int main() {
  std::mt19937_64 g(1);
  double rel_log = 0, abs_near1 = 0, abs_sc = 0, rel_sq = 0;
  for (long i = 0; i < 1000000; i++) {
    uint64_t k = g() >> 11, n = k + 1;
    if (i < 64) n = 1ull << (i %% 54);                      // exact powers of two, incl. n = 2^53 (u = 1)
    else if (i < 128) n = (1ull << 53) - (uint64_t)(i - 63);  // u just below 1
    else if (i < 4000) n = (g() >> (11 + (i %% 53))) + 1;     // small u
    const double u = (double)n * 1.1102230246251565404e-16;
    const double a = fast_log_u53(n, kLogTabRaw), b = log(u);
    if (fabs(b) > 1e-12) rel_log = fmax(rel_log, fabs(a - b) / fabs(b)); else abs_near1 = fmax(abs_near1, fabs(a - b));
    const double x = -2 * b, sq = fast_sqrt_pos(x), sr = sqrt(x);
    if (sr > 0) rel_sq = fmax(rel_sq, fabs(sq - sr) / sr); else if (sq != 0) rel_sq = 1;
    double s, c;
    fast_sincos_tab_u53(k, kLogTabRaw + MCMC_SCTAB_OFFSET, &s, &c);
    const long double th = 2.0L * 3.14159265358979323846264338327950288L * (long double)k / 9007199254740992.0L;
    abs_sc = fmax(abs_sc, fmax(fabs(s - (double)sinl(th)), fabs(c - (double)cosl(th))));
  }
  printf("%%.3e %%.3e %%.3e %%.3e\n", rel_log, abs_near1, rel_sq, abs_sc);
  return 0;
}
'''


def test_fastmath_matches_libm(tmp_path):
    src = tmp_path / "fm.cpp"
    src.write_text(SRC % ROOT)
    exe = tmp_path / "fm"
    subprocess.check_call(["/usr/bin/g++", "-O2", "-std=c++17", "-ffp-contract=off", str(src), "-o", str(exe)])
    rel_log, abs_near1, rel_sq, abs_sc = [float(x) for x in subprocess.check_output([str(exe)]).split()]
    assert rel_log <= 4.5e-16      # <= 2 ulp
    assert abs_near1 <= 3e-17      # log(u) for u within 1e-12 of 1
    assert rel_sq <= 2.3e-16       # <= 1 ulp
    assert abs_sc <= 1.7e-16


SRC_EXP = r'''
#include <cstdio>
#include <cmath>
#include <cstdint>
#include <random>
#include "%s/mcmc_b200/csrc/fastmath.cuh"
using namespace mcmc;
int main() {
  std::mt19937_64 g(2);
  std::uniform_real_distribution<double> U(0.0, 1.0);
  double rel = 0;
  for (long i = 0; i < 2000000; i++) {
    double x = (i %% 3 == 0) ? -U(g) * 700.0 : ((i %% 3 == 1) ? -U(g) * 5.0 : -exp(-40.0 * U(g)));
    if (i < 8) x = -0.25 * (double)i;
    const double a = exp_nonpos(x);
    const long double b = expl((long double)x);
    rel = fmax(rel, (double)(fabsl((long double)a - b) / b));
  }
  const double z = exp_nonpos(0.0), tiny = exp_nonpos(-800.0), nn = exp_nonpos(NAN), mi = exp_nonpos(-INFINITY);
  printf("%%.3e %%.17g %%.17g %%d %%.17g\n", rel, z, tiny, (int)(nn != nn), mi);
  return 0;
}
'''


def test_exp_nonpos_matches_libm(tmp_path):
    src = tmp_path / "fe.cpp"
    src.write_text(SRC_EXP % ROOT)
    exe = tmp_path / "fe"
    subprocess.check_call(["/usr/bin/g++", "-O2", "-std=c++17", "-ffp-contract=off", str(src), "-o", str(exe)])
    rel, z, tiny, isnan, mi = subprocess.check_output([str(exe)]).split()
    assert float(rel) <= 2.3e-16       # <= 1 ulp against long double
    assert float(z) == 1.0 and float(tiny) == 0.0 and int(isnan) == 1 and float(mi) == 0.0
