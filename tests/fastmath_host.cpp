// Host check of grackle_b200/csrc/fastmath.cuh against libm (long double) over the argument
// ranges of the solve path.  Prints "name max_ulp max_rel" per function; run by
// tests/test_fastmath_host.py.
#include <cstdio>
#include <cmath>
#include <cstdint>
#include <random>
#include "../grackle_b200/csrc/fastmath.cuh"

static double ulp_of(double x) { return std::fabs(std::nextafter(x, INFINITY) - x); }

template <class F, class R>
static void sweep(const char *name, F f, R ref, double lo, double hi, bool logspace, int n)
{
  std::mt19937_64 g(12345);
  std::uniform_real_distribution<double> u(0.0, 1.0);
  double worst_ulp = 0, worst_rel = 0, at = 0;
  for (int i = 0; i < n; i++) {
    double x = logspace ? std::exp(std::log(lo) + (std::log(hi) - std::log(lo)) * u(g)) : lo + (hi - lo) * u(g);
    const double got = f(x);
    const long double want = ref((long double)x);
    const double err = (double)fabsl((long double)got - want);
    const double ul = err / ulp_of((double)want);
    const double rel = want != 0 ? err / (double)fabsl(want) : err;
    if (ul > worst_ulp) { worst_ulp = ul; at = x; }
    if (rel > worst_rel) worst_rel = rel;
  }
  std::printf("%s %.3f %.3e %.17g\n", name, worst_ulp, worst_rel, at);
}

template <class F, class R>
static void sweep_abs(const char *name, F f, R ref, double lo, double hi, bool logspace, int n)
{
  std::mt19937_64 g(4321);
  std::uniform_real_distribution<double> u(0.0, 1.0);
  double worst = 0, at = 0;
  for (int i = 0; i < n; i++) {
    double x = logspace ? std::exp(std::log(lo) + (std::log(hi) - std::log(lo)) * u(g)) : lo + (hi - lo) * u(g);
    const long double want = ref((long double)x);
    const double err = (double)(fabsl((long double)f(x) - want) / fmaxl(1.0L, fabsl(want)));
    if (err > worst) { worst = err; at = x; }
  }
  std::printf("%s %.3e %.17g\n", name, worst, at);
}

int main()
{
  const int N = 2000000;
  static double tab[GRB_FM_TABLE_DOUBLES];
  grb::fm_fill_tables(tab);
  // table-driven versions: error relative to max(1, |result|) for log, relative for exp
  sweep_abs("tlog_wide", [&](double x) { return grb::dlog(x, tab); }, [](long double x) { return logl(x); }, 1e-300, 1e300, true, N);
  sweep_abs("tlog_near1", [&](double x) { return grb::dlog(x, tab); }, [](long double x) { return logl(x); }, 0.25, 4.0, false, N);
  sweep("tlog_far", [&](double x) { return grb::dlog(x, tab); }, [](long double x) { return logl(x); }, 3.0, 1e12, true, N);
  sweep("texp", [&](double x) { return grb::dexp(x, tab); }, [](long double x) { return expl(x); }, -700.0, 700.0, false, N);
  sweep("texp_small", [&](double x) { return grb::dexp(x, tab); }, [](long double x) { return expl(x); }, -2.0, 2.0, false, N);
  sweep("texp10", [&](double x) { return grb::dexp10(x, tab); }, [](long double x) { return powl(10.0L, x); }, -300.0, 300.0, false, N);
  sweep("tpow", [&](double x) { return grb::dpow_pos(x, 1.37, tab); }, [](long double x) { return powl(x, 1.37L); }, 1e-12, 1e12, true, N);
  sweep("log_wide", [](double x) { return grb::dlog(x); }, [](long double x) { return logl(x); }, 1e-300, 1e300, true, N);
  sweep("log_near1", [](double x) { return grb::dlog(x); }, [](long double x) { return logl(x); }, 0.5, 2.0, false, N);
  sweep("log10", [](double x) { return grb::dlog10(x); }, [](long double x) { return log10l(x); }, 1e-12, 1e12, true, N);
  sweep("exp", [](double x) { return grb::dexp(x); }, [](long double x) { return expl(x); }, -700.0, 700.0, false, N);
  sweep("exp_small", [](double x) { return grb::dexp(x); }, [](long double x) { return expl(x); }, -2.0, 2.0, false, N);
  sweep("exp10", [](double x) { return grb::dexp10(x); }, [](long double x) { return powl(10.0L, x); }, -300.0, 300.0, false, N);
  sweep("exp10_small", [](double x) { return grb::dexp10(x); }, [](long double x) { return powl(10.0L, x); }, -40.0, 5.0, false, N);
  sweep("rcp", [](double x) { return grb::drcp(x); }, [](long double x) { return 1.0L / x; }, 1e-30, 1e30, true, N);
  sweep("div", [](double x) { return grb::ddiv(0.7310585786300049, x); }, [](long double x) { return 0.7310585786300049L / x; }, 1e-30, 1e30, true, N);
  sweep("pow", [](double x) { return grb::dpow_pos(x, 1.37); }, [](long double x) { return powl(x, 1.37L); }, 1e-12, 1e12, true, N);
  // special arguments must behave like libm
  const double sp[] = {0.0, -1.0, INFINITY, NAN, 1e-310};
  for (double x : sp) {
    const double a = grb::dlog(x), b = std::log(x);
    if (!((a == b) || (std::isnan(a) && std::isnan(b)))) { std::printf("special_log FAIL %g\n", x); return 1; }
  }
  const double se[] = {-1e4, 1e4, -745.0, 709.5, INFINITY, -INFINITY, NAN};
  for (double x : se) {
    const double a = grb::dexp(x), b = std::exp(x);
    if (!((a == b) || (std::isnan(a) && std::isnan(b)))) { std::printf("special_exp FAIL %g\n", x); return 1; }
  }
  std::printf("special ok\n");
  return 0;
}
