// grackle-b200: host-side generation of the temperature-tabulated reaction and
// cooling coefficients (600 ln-T bins by default) that the device kernels
// interpolate.  Runs once in _initialize_chemistry_data.
//
// What is tabulated, and the analytic fits used, follow the reference's
// initialize_rates.c:213-472 (which tables, which unit factor) and
// rate_functions.c:32-1339 (the fits: Abel+97 k1-k19, Glover&Abel08 H2 cooling,
// Galli&Palla99, Lique15, Coppola+11 HD, Ripamonti&Abel04 CIE, ...).  The code
// here is organised differently (one descriptor table, shared polynomial
// helpers) but evaluates every fit with the same floating-point operation
// order so the tables agree with the reference's to the last bit; tests/
// check that against oracle/_ref (the reference's own C compiled unmodified).
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <algorithm>

#include "grackle.h"
#include "gr_internal.h"

namespace {

constexpr double kTiny = 1.0e-20;            // grackle_macros.h:113
constexpr double kHugeExpArg = 1.0e30;       // "dhuge", rate_functions.c:18
constexpr double kBoltz = 1.3806504e-16;     // phys_constants.h
constexpr double kMh = 1.67262171e-24;
constexpr double kTevK = 1.1605e4;           // rate_functions.c:14

// c0 + c1*x + c2*x^2 + ... summed left to right with pow() for the powers,
// which is how the reference spells its fits.
template <int N>
inline double series(double x, const double (&c)[N])
{
  double s = c[0];
  if (N > 1) s += c[1] * x;
  // pow(x, 2) with a literal 2 is folded to x*x by gcc (the compiler the reference and the
  // oracle are built with); glibc's pow(x, 2.0) is not always the correctly rounded square
  if (N > 2) s += c[2] * (x * x);
  for (int n = 3; n < N; ++n) s += c[n] * std::pow(x, (double)n);
  return s;
}

inline double expcap(double arg) { return std::exp(-std::fmin(std::log(kHugeExpArg), arg)); }

struct Ctx { const chemistry_data *p; };

// ---------------------------------------------------------------------------
// chemistry rate coefficients (cgs, divided by `u`)
// ---------------------------------------------------------------------------
double r_k1(double T, double u, const Ctx &)
{
  static const double c[] = {-32.71396786375, 13.53655609057, -5.739328757388,
                             1.563154982022, -0.2877056004391, 0.03482559773736999,
                             -0.00263197617559, 0.0001119543953861, -2.039149852002e-6};
  const double tev = T / 11605.0;
  double k = std::exp(series(std::log(tev), c)) / u;
  if (tev <= 0.8) k = std::fmax(kTiny, k);
  return k;
}

double r_k3(double T, double u, const Ctx &)
{
  static const double c[] = {-44.09864886561001, 23.91596563469, -10.75323019821,
                             3.058038757198, -0.5685118909884001, 0.06795391233790001,
                             -0.005009056101857001, 0.0002067236157507, -3.649161410833e-6};
  const double tev = T / 11605.0;
  if (tev > 0.8) return std::exp(series(std::log(tev), c)) / u;
  return kTiny;
}

double r_k4(double T, double u, const Ctx &cx)
{
  const double tev = T / 11605.0;
  if (cx.p->CaseBRecombination == 1) return 1.26e-14 * std::pow(5.7067e5 / T, 0.75) / u;
  if (tev > 0.8)
    return (1.54e-9 * (1.0 + 0.3 / std::exp(8.099328789667 / tev)) /
                (std::exp(40.49664394833662 / tev) * std::pow(tev, 1.5)) +
            3.92e-13 / std::pow(tev, 0.6353)) / u;
  return 3.92e-13 / std::pow(tev, 0.6353) / u;
}

double r_k2(double T, double u, const Ctx &cx)
{
  if (cx.p->CaseBRecombination == 1) {
    if (T < 1.0e9)
      return 4.881357e-6 * std::pow(T, -1.5) *
             std::pow((1.0 + 1.14813e2 * std::pow(T, -0.407)), -2.242) / u;
    return kTiny;
  }
  if (T > 5500) {
    static const double c[] = {-28.61303380689232, -0.7241125657826851, -0.02026044731984691,
                               -0.002380861877349834, -0.0003212605213188796,
                               -0.00001421502914054107, 4.989108920299513e-6,
                               5.755614137575758e-7, -1.856767039775261e-8,
                               -3.071135243196595e-9};
    return std::exp(series(std::log(T / kTevK), c)) / u;
  }
  return r_k4(T, u, cx);
}

double r_k5(double T, double u, const Ctx &)
{
  static const double c[] = {-68.71040990212001, 43.93347632635, -18.48066993568,
                             4.701626486759002, -0.7692466334492, 0.08113042097303,
                             -0.005324020628287001, 0.0001975705312221, -3.165581065665e-6};
  const double tev = T / 11605.0;
  if (tev > 0.8) return std::exp(series(std::log(tev), c)) / u;
  return kTiny;
}

double r_k6(double T, double u, const Ctx &cx)
{
  if (cx.p->CaseBRecombination == 1) {
    if (T < 1.0e9)
      return 7.8155e-5 * std::pow(T, -1.5) *
             std::pow((1.0 + 2.0189e2 * std::pow(T, -0.407)), -2.242) / u;
    return kTiny;
  }
  return 3.36e-10 / std::sqrt(T) / std::pow(T / 1.0e3, 0.2) / (1.0 + std::pow(T / 1.0e6, 0.7)) / u;
}

double r_k7(double T, double u, const Ctx &)
{
  return 3.0e-16 * std::pow(T / 3.0e2, 0.95) * std::exp(-T / 9.32e3) / u;
}

double r_k8(double T, double u, const Ctx &)
{
  return 1.35e-9 *
         (std::pow(T, 9.8493e-2) + 3.2852e-1 * std::pow(T, 5.5610e-1) + 2.771e-7 * std::pow(T, 2.1826)) /
         (1.0 + 6.191e-3 * std::pow(T, 1.0461) + 8.9712e-11 * std::pow(T, 3.0424) +
          3.2576e-14 * std::pow(T, 3.7741)) / u;
}

double r_k9(double T, double u, const Ctx &)
{
  if (T < 30.0) return 2.10e-20 * std::pow(T / 30.0, -0.15) / u;
  const double l = std::log10(std::fmin(T, 3.2e4));
  return std::pow(10.0, -18.20 - 3.194 * l + 1.786 * std::pow(l, 2) - 0.2072 * std::pow(l, 3.0)) / u;
}

double r_k10(double, double u, const Ctx &) { return 6.0e-10 / u; }

double r_k11(double T, double u, const Ctx &cx)
{
  const double tev = T / 11605.0;
  if (!(tev > 0.3)) return kTiny;
  if (cx.p->h2_charge_exchange_rate == 1) {
    static const double c[] = {-3.3232183e-07, 3.3735382e-07, -1.4491368e-07, 3.4172805e-08,
                               -4.7813720e-09, 3.9731542e-10, -1.8171411e-11, 3.5311932e-13};
    return (std::exp(-21237.15 / T) * series(std::log(T), c)) / u;
  }
  if (cx.p->h2_charge_exchange_rate == 2) {
    static const double c[] = {-24.24914687731536, 3.400824447095291, -3.898003964650152,
                               2.045587822403071, -0.5416182856220388, 0.0841077503763412,
                               -0.007879026154483455, 0.0004138398421504563,
                               -9.36345888928611e-6};
    return std::exp(series(std::log(tev), c)) / u;
  }
  std::fprintf(stderr, "grackle-b200: h2_charge_exchange_rate must be 1 or 2 (got %d)\n",
               cx.p->h2_charge_exchange_rate);
  std::exit(0);
}

double r_k12(double T, double u, const Ctx &)
{
  if (T / 11605.0 > 0.3) return 4.4886e-9 * std::pow(T, 0.109127) * std::exp(-101858.0 / T) / u;
  return kTiny;
}

double r_k13(double T, double u, const Ctx &cx)
{
  const double tev = T / 11605.0;
  double k;
  switch (cx.p->three_body_rate) {
    case 0:
      if (tev > 0.3)
        k = 1.0670825e-10 * std::pow(tev, 2.012) /
            (std::exp(4.463 / tev) * std::pow((1.0 + 0.2472 * tev), 3.512));
      else
        k = kTiny * u;
      break;
    case 1: k = (5.24e-7 / std::pow(T, 0.485)) * std::exp(-5.2e4 / T); break;
    case 2: k = 8.4e-11 * std::pow(T, 0.515) * std::exp(-5.2e4 / T); break;
    case 3: k = (1.38e-4 / std::pow(T, 1.025)) * std::exp(-5.2e4 / T); break;
    case 4: {
      const double l = std::log10(T);
      k = std::pow(1e1, (-178.4239 - 68.42243 * l + 43.20243 * std::pow(l, 2.0) -
                         4.633167 * std::pow(l, 3.0) + 69.70086 * std::log10(1.0 + 40870.38 / T) -
                         (23705.7 / T)));
      break;
    }
    case 5:
      k = (T <= 3000.0) ? 2.4e-8 * std::exp(-5.2e4 / T)
                        : 2.2e-6 * std::pow(T, -0.565) * std::exp(-5.2e4 / T);
      break;
    default:
      std::fprintf(stderr, "grackle-b200: unknown three_body_rate %d\n", cx.p->three_body_rate);
      std::exit(0);
  }
  return k / u;
}

// Martin, Schwarz & Mandy (1996) density-dependent H2 collisional dissociation:
// 7 fit functions for direct dissociation (set 0) and 7 for tunnelling (set 1).
void r_k13dd(double T, double u, double out[14])
{
  static const double fit[2][21] = {
      {-1.784239e2, -6.842243e1, 4.320243e1, -4.633167e0, 6.970086e1, 4.087038e4, -2.370570e4,
       1.288953e2, -5.391334e1, 5.315517e0, -1.973427e1, 1.678095e4, -2.578611e4, 1.482123e1,
       -4.890915e0, 4.749030e-1, -1.338283e2, -1.164408e0, 8.227443e-1, 5.864073e-1, -2.056313e0},
      {-1.427664e+02, 4.270741e+01, -2.027365e+00, -2.582097e-01, 2.136094e+01, 2.753531e+04,
       -2.146779e+04, 6.034928e+01, -2.743096e+01, 2.676150e+00, -1.128215e+01, 1.425455e+04,
       -2.312520e+04, 9.305564e+00, -2.464009e+00, 1.985955e-01, 7.430600e+02, -1.174242e+00,
       7.502286e-01, 2.358848e-01, 2.937507e+00}};
  if (T <= 500.0) T = 500.0;
  if (T >= 1.0e6) T = 1.0e6;
  const double l = std::log10(T);
  for (int s = 0; s < 2; ++s) {
    const double *f = fit[s];
    const double a = f[0] + f[1] * l + f[2] * std::pow(l, 2) + f[3] * std::pow(l, 3.0) +
                     f[4] * std::log10(1.0 + f[5] / T);
    const double a1 = f[6] / T;
    const double b = f[7] + f[8] * l + f[9] * std::pow(l, 2.0) + f[10] * std::log10(1.0 + f[11] / T);
    const double b1 = f[12] / T;
    const double c = f[13] + f[14] * l + f[15] * std::pow(l, 2.0) + f[16] / T;
    const double c1 = f[17] + c;
    const double d = f[18] + f[19] * std::exp(-T / 1850.0) + f[20] * std::exp(-T / 440.0);
    double *o = out + 7 * s;
    o[0] = a - std::log10(u);
    o[1] = a - b;
    o[2] = a1;
    o[3] = a1 - b1;
    o[4] = std::pow(10.0, c);
    o[5] = std::pow(10.0, c1);
    o[6] = d;
  }
}

double r_k14(double T, double u, const Ctx &)
{
  static const double c[] = {-18.01849334273, 2.360852208681, -0.2827443061704, 0.01623316639567,
                             -0.03365012031362999, 0.01178329782711, -0.001656194699504,
                             0.0001068275202678, -2.631285809207e-6};
  const double tev = T / 11605.0;
  if (tev > 0.04) return std::exp(series(std::log(tev), c)) / u;
  return kTiny;
}

double r_k15(double T, double u, const Ctx &)
{
  static const double c[] = {-20.37260896533324, 1.139449335841631, -0.1421013521554148,
                             0.00846445538663, -0.0014327641212992, 0.0002012250284791,
                             0.0000866396324309, -0.00002585009680264, 2.4555011970392e-6,
                             -8.06838246118e-8};
  const double tev = T / 11605.0;
  if (tev > 0.1) return std::exp(series(std::log(tev), c)) / u;
  return 2.56e-9 * std::pow(tev, 1.78186) / u;
}

double r_k16(double T, double u, const Ctx &) { return 2.4e-6 * (1.0 + T / 2.0e4) / std::sqrt(T) / u; }

double r_k17(double T, double u, const Ctx &)
{
  if (T > 1.0e4) return 4.0e-4 * std::pow(T, -1.4) * std::exp(-15100.0 / T) / u;
  return 1.0e-8 * std::pow(T, -0.4) / u;
}

double r_k18(double T, double u, const Ctx &)
{
  if (T > 617.0) return 1.32e-6 * std::pow(T, -0.76) / u;
  return 1.0e-8 / u;
}

double r_k19(double T, double u, const Ctx &) { return 5.e-7 * std::sqrt(100.0 / T) / u; }
double r_k20(double, double, const Ctx &) { return kTiny; }
double r_k21(double T, double u, const Ctx &) { return 2.8e-31 * std::pow(T, -0.6) / u; }

double r_k22(double T, double u, const Ctx &cx)
{
  double k;
  switch (cx.p->three_body_rate) {
    case 0:
      k = (T <= 300.0) ? 1.3e-32 * std::pow(T / 300.0, -0.38) : 1.3e-32 * std::pow(T / 300.0, -1.0);
      break;
    case 1: k = 5.5e-29 / T; break;
    case 2: k = 8.8e-33; break;
    case 3: k = 1.44e-26 / std::pow(T, 1.54); break;
    case 4: k = 7.7e-31 / std::pow(T, 0.464); break;
    case 5: k = (6e-32 / std::pow(T, 0.25)) + (2e-31 / std::pow(T, 0.5)); break;
    default:
      std::fprintf(stderr, "grackle-b200: unknown three_body_rate %d\n", cx.p->three_body_rate);
      std::exit(0);
  }
  return k / u;
}

double r_k23(double T, double u, const Ctx &)
{
  const double k = ((8.125e-8 / std::sqrt(T)) * std::exp(-52000.0 / T) * (1.0 - std::exp(-6000.0 / T))) / u;
  return (kTiny > k) ? kTiny : k;
}

double r_k50(double T, double u, const Ctx &)
{
  if (T <= 2.0e5)
    return (2.0e-10 * std::pow(T, 0.402) * std::exp(-3.71e1 / T) - 3.31e-17 * std::pow(T, 1.48)) / u;
  return 2.5e-8 * std::pow(T / 2.0e5, 0.402) / u;
}

double r_k51(double T, double u, const Ctx &)
{
  return (2.06e-10 * std::pow(T, 0.396) * std::exp(-3.30e1 / T) + 2.03e-9 * std::pow(T, -0.332)) / u;
}

double r_k52(double T, double u, const Ctx &)
{
  if (T <= 1e4)
    return 1.0e-9 * (0.417 + 0.846 * std::log10(T) - 0.137 * std::pow(std::log10(T), 2.0)) / u;
  return 1.609e-9 / u;
}

double r_k53(double T, double u, const Ctx &) { return 1.1e-9 * std::exp(-4.88e2 / T) / u; }

// NB: the reference does not divide k54 by the unit factor (rate_functions.c:627-638)
double r_k54(double T, double, const Ctx &)
{
  if (T <= 2.0e3) {
    static const double c[] = {-5.64737e1, 5.88886, 7.19692, 2.25069, -2.16903, 3.17887e-1};
    return std::pow(1.0e1, series(std::log10(T), c));
  }
  return 3.17e-10 * std::exp(-5.207e3 / T);
}

double r_k55(double T, double u, const Ctx &)
{
  if (T <= 2.0e2) return 1.08e-22 / u;
  return 5.25e-11 * std::exp(-4.43e3 / T + 1.739e5 / std::pow(T, 2.0)) / u;
}

double r_k57(double T, double u, const Ctx &)
{
  if (T > 3.0e3) return 1.2e-17 * std::pow(T, 1.2) * std::exp(-1.578e5 / T) / u;
  return kTiny;
}

double r_k58(double T, double u, const Ctx &)
{
  if (T > 3.0e3) return 1.75e-17 * std::pow(T, 1.3) * std::exp(-1.578e5 / T) / u;
  return kTiny;
}

double r_h2dust(double T, double Td, double u, const Ctx &cx)
{
  const double fgr = 0.009387;
  double r;
  if (cx.p->h2_dust_rate == 1) {
    r = 6.0e-17 / fgr * std::pow(T / 300.0, 0.5) *
        (std::pow(1.0 + std::exp(7.5e2 * ((1.0 / 75.0) - (1.0 / Td))), -1.0)) *
        (std::pow(1.0 + (4.0e-2 * std::pow(T + Td, 0.5)) + (2.0e-3 * T) + (8.0e-6 * std::pow(T, 2.0)), -1.0));
  } else {
    const double t2 = T / 1.0e2, td2 = Td / 1.0e2;
    r = 3.0e-17 / fgr * std::pow(t2, 0.5) /
        (1.0 + 0.4 * std::pow(t2 + td2, 0.5) + 0.2 * t2 + 8.0e-2 * std::pow(t2, 2.0));
  }
  return r / u;
}

// critical-density pieces for H2 formation heating (Omukai 2000); not unit-scaled
double r_ncr_n(double T, double, const Ctx &) { return 1.0e6 * std::pow(T, -0.5); }
double r_ncr_d1(double T, double, const Ctx &) { return 1.6 * std::exp(-std::pow(400.0 / T, 2.0)); }
double r_ncr_d2(double T, double, const Ctx &) { return 1.4 * std::exp(-12000.0 / (T + 1200.0)); }

// ---------------------------------------------------------------------------
// cooling coefficients (erg cm^3 s^-1 style, divided by `u`)
// ---------------------------------------------------------------------------
double c_ceHI(double T, double u, const Ctx &cx)
{
  if (cx.p->collisional_excitation_rates != 1) return kTiny;
  return 7.5e-19 * expcap(118348.0 / T) / (1.0 + std::sqrt(T / 1.0e5)) / u;
}
double c_ceHeI(double T, double u, const Ctx &cx)
{
  if (cx.p->collisional_excitation_rates != 1) return kTiny;
  return 9.1e-27 * expcap(13179.0 / T) * std::pow(T, -0.1687) / (1.0 + std::sqrt(T / 1.0e5)) / u;
}
double c_ceHeII(double T, double u, const Ctx &cx)
{
  if (cx.p->collisional_excitation_rates != 1) return kTiny;
  return 5.54e-17 * expcap(473638.0 / T) * std::pow(T, -0.3970) / (1.0 + std::sqrt(T / 1.0e5)) / u;
}
double c_ciHeIS(double T, double u, const Ctx &cx)
{
  if (cx.p->collisional_ionisation_rates != 1) return kTiny;
  return 5.01e-27 * std::pow(T, -0.1687) / (1.0 + std::sqrt(T / 1.0e5)) * expcap(55338.0 / T) / u;
}
double c_ciHI(double T, double u, const Ctx &cx)
{
  if (cx.p->collisional_ionisation_rates != 1) return kTiny;
  return 2.18e-11 * r_k1(T, 1, cx) / u;
}
double c_ciHeI(double T, double u, const Ctx &cx)
{
  if (cx.p->collisional_ionisation_rates != 1) return kTiny;
  return 3.94e-11 * r_k3(T, 1, cx) / u;
}
double c_ciHeII(double T, double u, const Ctx &cx)
{
  if (cx.p->collisional_ionisation_rates != 1) return kTiny;
  return 8.72e-11 * r_k5(T, 1, cx) / u;
}

// Hui & Gnedin (1997) style recombination cooling, case A or B
double recomb_hg97(double T, double lam, bool caseB)
{
  if (caseB)
    return 3.435e-30 * T * std::pow(lam, 1.970) / std::pow(1.0 + std::pow(lam / 2.25, 0.376), 3.720);
  return 1.778e-29 * T * std::pow(lam, 1.965) / std::pow(1.0 + std::pow(lam / 0.541, 0.502), 2.697);
}
double c_reHII(double T, double u, const Ctx &cx)
{
  if (cx.p->recombination_cooling_rates != 1) return kTiny;
  return recomb_hg97(T, 2.0 * 157807.0 / T, cx.p->CaseBRecombination == 1) / u;
}
double c_reHeII1(double T, double u, const Ctx &cx)
{
  if (cx.p->recombination_cooling_rates != 1) return kTiny;
  const double lam = 2.0 * 285335.0 / T;
  if (cx.p->CaseBRecombination == 1) return 1.26e-14 * kBoltz * T * std::pow(lam, 0.75) / u;
  return 3e-14 * kBoltz * T * std::pow(lam, 0.654) / u;
}
double c_reHeII2(double T, double u, const Ctx &cx)
{
  if (cx.p->recombination_cooling_rates != 1) return kTiny;
  return 1.24e-13 * std::pow(T, -1.5) * expcap(470000.0 / T) * (1.0 + 0.3 * expcap(94000.0 / T)) / u;
}
double c_reHeIII(double T, double u, const Ctx &cx)
{
  if (cx.p->recombination_cooling_rates != 1) return kTiny;
  const double lam = 2.0 * 631515.0 / T;
  if (cx.p->CaseBRecombination == 1)
    return 8.0 * 3.435e-30 * T * std::pow(lam, 1.970) /
           std::pow(1.0 + std::pow(lam / 2.25, 0.376), 3.720) / u;
  return 8.0 * 1.778e-29 * T * std::pow(lam, 1.965) /
         std::pow(1.0 + std::pow(lam / 0.541, 0.502), 2.697) / u;
}
double c_brem(double T, double u, const Ctx &cx)
{
  if (cx.p->bremsstrahlung_cooling_rates != 1) return kTiny;
  return 1.43e-27 * std::sqrt(T) * (1.1 + 0.34 * std::exp(-std::pow(5.5 - std::log10(T), 2.0) / 3.0)) / u;
}

// Lepp & Shull (1983) H2 cooling pieces (tabulated for ABI parity; the kernels use
// the Glover & Abel (2008) branch like the reference does)
double ls83_par(double T)
{
  if (T > 1635.0) return 1.0e-12 * std::sqrt(T) * std::exp(-1000.0 / T);
  return 1.4e-13 * std::exp((T / 125.0) - std::pow(T / 577.0, 2.0));
}
double c_vibh(double T, double u, const Ctx &) { return 1.1e-18 * expcap(6744.0 / T) / u; }
double c_hyd01k(double T, double u, const Ctx &)
{
  return ls83_par(T) * expcap(8.152e-13 / (kBoltz * T)) / u;
}
double c_h2k01(double T, double u, const Ctx &)
{
  const double arg = 8.152e-13 * (4.2 / (kBoltz * (T + 1190.0)) + 1.0 / (kBoltz * T));
  return 1.45e-12 * std::sqrt(T) * expcap(arg) / u;
}
double c_rotl(double T, double u, const Ctx &)
{
  const double x = std::log10(T / 1.0e4);
  if (T > 4031.0) return 1.38e-22 * std::exp(-9243.0 / T) / u;
  return std::pow(10.0, -22.9 - 0.553 * x - 1.148 * std::pow(x, 2.0)) / u;
}
double c_roth(double T, double u, const Ctx &)
{
  const double x = std::log10(T / 1.0e4);
  if (T > 1087.0) return 3.9e-19 * std::exp(-6118.0 / T) / u;
  return std::pow(10.0, -19.24 + 0.474 * x - 1.247 * std::pow(x, 2.0)) / u;
}

// Galli & Palla (1998)
double c_gp99low(double T, double u, const Ctx &)
{
  const double lt = std::log10(std::fmin(std::fmax(T, 13.0), 1.0e5));
  return std::pow(10.0, -103.0 + 97.59 * lt - 48.05 * std::pow(lt, 2) + 10.8 * std::pow(lt, 3.0) -
                            0.9032 * std::pow(lt, 4.0)) / u;
}
double c_gp99high(double T, double u, const Ctx &)
{
  const double t3 = std::fmin(std::fmax(T, 13.0), 1.0e5) / 1000.0;
  const double lr = (9.5e-22 * std::pow(t3, 3.76)) / (1.0 + 0.12 * std::pow(t3, 2.1)) *
                        std::exp(-std::pow((0.13 / t3), 3.0)) + 3.0e-24 * std::exp(-0.51 / t3);
  const double lv = 6.7e-19 * std::exp(-5.86 / t3) + 1.6e-18 * std::exp(-11.7 / t3);
  return (lr + lv) / u;
}

// Glover & Abel (2008) low-density H2 cooling by collision partner
inline double ga_lt3(double T, double &tm)
{
  tm = std::fmin(std::fmax(T, 10.0), 1.0e4);
  return std::log10(tm / 1.0e3);
}
double c_gaHI(double T, double u, const Ctx &cx)
{
  double tm;
  const double l = ga_lt3(T, tm);
  if (cx.p->h2_h_cooling_rate == 1) {   // Lique (2015)
    if (tm < 1e2) return 0.0;
    static const double c[] = {-24.07950609, 4.54182810, -2.40206896, -0.75355292,
                               4.69258178, -2.79573574, -3.14766075, 2.50751333};
    return std::pow(10.0, series(l, c)) / u;
  }
  if (cx.p->h2_h_cooling_rate == 2) {
    static const double lo[] = {-16.818342, 37.383713, 58.145166, 48.656103, 20.159831, 3.8479610};
    static const double mid[] = {-24.311209, 3.5692468, -11.332860, -27.850082, -21.328264, -4.2519023};
    static const double hi[] = {-24.311209, 4.6450521, -3.7209846, 5.9369081, -5.5108047, 1.5538288};
    if (tm < 1.0e2) return std::pow(10.0, series(l, lo)) / u;
    if (tm < 1.0e3) return std::pow(10.0, series(l, mid)) / u;
    return std::pow(10.0, series(l, hi)) / u;
  }
  std::fprintf(stderr, "grackle-b200: h2_h_cooling_rate must be 1 or 2 (got %d)\n",
               cx.p->h2_h_cooling_rate);
  std::exit(0);
}
double c_gaH2(double T, double u, const Ctx &)
{
  static const double c[] = {-23.962112, 2.09433740, -0.77151436, 0.43693353, -0.14913216, -0.033638326};
  double tm;
  return std::pow(10.0, series(ga_lt3(T, tm), c)) / u;
}
double c_gaHe(double T, double u, const Ctx &)
{
  static const double c[] = {-23.689237, 2.1892372, -0.81520438, 0.29036281, -0.16596184, 0.19191375};
  double tm;
  return std::pow(10.0, series(ga_lt3(T, tm), c)) / u;
}
double c_gaHp(double T, double u, const Ctx &)
{
  static const double c[] = {-22.089523, 1.5714711, 0.015391166, -0.23619985, -0.51002221, 0.32168730};
  double tm;
  return std::pow(10.0, series(ga_lt3(T, tm), c)) / u;
}
double c_gael(double T, double u, const Ctx &)
{
  static const double lo[] = {-21.928796, 16.815730, 96.743155, 343.19180, 734.71651,
                              983.67576, 801.81247, 364.14446, 70.609154};
  static const double hi[] = {-22.921189, 1.6802758, 0.93310622, 4.0406627, -4.7274036,
                              -8.8077017, 8.9167183, 6.4380698, -6.3701156};
  double tm;
  const double l = ga_lt3(T, tm);
  if (tm < 100.0) return 0.0;
  if (tm < 500.0) return std::pow(10.0, series(l, lo)) / u;
  return std::pow(10.0, series(l, hi)) / u;
}
double c_h2lte(double T, double u, const Ctx &)
{
  static const double c[] = {-20.584225, 5.0194035, -1.5738805, -4.7155769, 2.4714161,
                             5.4710750, -3.9467356, -2.2148338, 1.8161874};
  double tm;
  const double l = ga_lt3(T, tm);
  if (tm < 1.0e2) return 7.0e-27 * std::pow(tm, 1.5) * std::exp(-512.0 / tm) / u;
  return std::pow(10.0, series(l, c)) / u;
}

// HD cooling (Coppola+11 LTE, Wrathmall+07 low density)
double c_hdlte(double T, double u, const Ctx &)
{
  static const double c[] = {-55.5725, 56.649, -37.9102, 12.698, -2.02424, 0.122393};
  const double l = std::log10(std::fmin(std::fmax(T, 10.0), 3.0e4));
  return std::pow(10.0, std::fmin(series(l, c), 0.0)) / u;
}
double c_hdlow(double T, double u, const Ctx &)
{
  static const double c[] = {-23.175780, 1.5035261, 0.40871403, 0.17849311, -0.077291388, 0.10031326};
  const double l = std::log10(std::fmin(std::fmax(T, 1.0e1), 6.0e3) / 1.0e3);
  return std::pow(10.0, series(l, c)) / u;
}

// H2-H2 / H2-He collision-induced emission, optically thin (Ripamonti & Abel 2004):
// piecewise-linear table in T with power-law tails.  Data: see cie_table_data.inc.
#include "cie_table_data.inc"
double cie_thin(double T)
{
  const int last = kCieCount - 1;
  if (T <= kCieT[0]) return kCieRate[0] * std::pow(T / kCieT[0], 4.0);
  if (T >= kCieT[last]) return kCieRate[last] * std::pow(T / kCieT[last], 3.0);
  int lo = 0, hi = last;
  while (hi - lo > 1) {           // same midpoint bisection as rate_functions.c:1259-1276
    const int mid = (hi + lo) / 2;
    if (T >= kCieT[mid]) lo = mid; else hi = mid;
  }
  return (kCieRate[hi] * (T - kCieT[lo]) + kCieRate[lo] * (kCieT[hi] - T)) / (kCieT[hi] - kCieT[lo]);
}
double c_cieco(double T, double u, const Ctx &) { return cie_thin(T) * (kMh / 2.0) / u; }

// dust: gas-grain energy transfer (Hollenbach & McKee 89) and grain recombination (Wolfire+95)
double c_gasgrain(double T, double u, const Ctx &)
{
  const double fgr = 0.009387;
  const double coeff = 1.2e-31 * std::pow(1.0e3, -0.5) / fgr;
  return coeff * std::pow(T, 0.5) * (1.0 - 0.8 * std::exp(-75.0 / T)) / u;
}
double c_regr(double T, double u, const Ctx &)
{
  const double beta = 0.74 / std::pow(T, 0.068);
  return 4.65e-30 * std::pow(T, 0.94 + 0.5 * beta) / u;
}

// ---------------------------------------------------------------------------
typedef double (*fit_fn)(double, double, const Ctx &);
enum UnitKind { U_K, U_K3, U_COOL };
struct TableSpec { double *chemistry_data_storage::*slot; fit_fn fn; UnitKind unit; };

#define SLOT(m) &chemistry_data_storage::m
const TableSpec kPrimordialTables[] = {
    {SLOT(k1), r_k1, U_K},     {SLOT(k2), r_k2, U_K},     {SLOT(k3), r_k3, U_K},
    {SLOT(k4), r_k4, U_K},     {SLOT(k5), r_k5, U_K},     {SLOT(k6), r_k6, U_K},
    {SLOT(k7), r_k7, U_K},     {SLOT(k8), r_k8, U_K},     {SLOT(k9), r_k9, U_K},
    {SLOT(k10), r_k10, U_K},   {SLOT(k11), r_k11, U_K},   {SLOT(k12), r_k12, U_K},
    {SLOT(k13), r_k13, U_K},   {SLOT(k14), r_k14, U_K},   {SLOT(k15), r_k15, U_K},
    {SLOT(k16), r_k16, U_K},   {SLOT(k17), r_k17, U_K},   {SLOT(k18), r_k18, U_K},
    {SLOT(k19), r_k19, U_K},   {SLOT(k20), r_k20, U_K},   {SLOT(k21), r_k21, U_K3},
    {SLOT(k22), r_k22, U_K3},  {SLOT(k23), r_k23, U_K},
    {SLOT(k50), r_k50, U_K},   {SLOT(k51), r_k51, U_K},   {SLOT(k52), r_k52, U_K},
    {SLOT(k53), r_k53, U_K},   {SLOT(k54), r_k54, U_K},   {SLOT(k55), r_k55, U_K},
    {SLOT(k56), r_k8, U_K},    {SLOT(k57), r_k57, U_K},   {SLOT(k58), r_k58, U_K},
    {SLOT(n_cr_n), r_ncr_n, U_K}, {SLOT(n_cr_d1), r_ncr_d1, U_K}, {SLOT(n_cr_d2), r_ncr_d2, U_K},
    {SLOT(ceHI), c_ceHI, U_COOL},       {SLOT(ceHeI), c_ceHeI, U_COOL},
    {SLOT(ceHeII), c_ceHeII, U_COOL},   {SLOT(ciHeIS), c_ciHeIS, U_COOL},
    {SLOT(ciHI), c_ciHI, U_COOL},       {SLOT(ciHeI), c_ciHeI, U_COOL},
    {SLOT(ciHeII), c_ciHeII, U_COOL},   {SLOT(reHII), c_reHII, U_COOL},
    {SLOT(reHeII1), c_reHeII1, U_COOL}, {SLOT(reHeII2), c_reHeII2, U_COOL},
    {SLOT(reHeIII), c_reHeIII, U_COOL}, {SLOT(brem), c_brem, U_COOL},
    {SLOT(vibh), c_vibh, U_COOL},       {SLOT(hyd01k), c_hyd01k, U_COOL},
    {SLOT(h2k01), c_h2k01, U_COOL},     {SLOT(rotl), c_rotl, U_COOL},
    {SLOT(roth), c_roth, U_COOL},
    {SLOT(GP99LowDensityLimit), c_gp99low, U_COOL},
    {SLOT(GP99HighDensityLimit), c_gp99high, U_COOL},
    {SLOT(GAHI), c_gaHI, U_COOL},       {SLOT(GAH2), c_gaH2, U_COOL},
    {SLOT(GAHe), c_gaHe, U_COOL},       {SLOT(GAHp), c_gaHp, U_COOL},
    {SLOT(GAel), c_gael, U_COOL},       {SLOT(H2LTE), c_h2lte, U_COOL},
    {SLOT(HDlte), c_hdlte, U_COOL},     {SLOT(HDlow), c_hdlow, U_COOL},
    {SLOT(cieco), c_cieco, U_COOL},
};
const TableSpec kDustTables[] = {
    {SLOT(gas_grain), c_gasgrain, U_COOL},
    {SLOT(regr), c_regr, U_COOL},
};
#undef SLOT

}  // namespace

// Fill every temperature table of `rates` for the unit system in `units`.
// Unit factors: initialize_rates.c:254-285.
int gr_build_rate_tables(chemistry_data *chem, chemistry_data_storage *rates, const code_units *units,
                         double co_length_units, double co_density_units)
{
  const Ctx cx{chem};
  const int nbin = chem->NumberOfTemperatureBins;
  const double aprod = units->a_value * units->a_units;
  const double tbase = units->time_units;
  const double xbase = co_length_units / aprod;
  const double dbase = co_density_units * std::pow(aprod, 3.0);
  const double u_k = (std::pow(units->a_units, 3.0) * kMh) / (dbase * tbase);
  const double u_k3 = u_k * (std::pow(units->a_units, 3.0) * kMh) / dbase;
  const double u_cool = (std::pow(units->a_units, 5) * std::pow(xbase, 2) * std::pow(kMh, 2.0)) /
                        (dbase * std::pow(tbase, 3.0));
  const double unit_of[3] = {u_k, u_k3, u_cool};

  const double lnT0 = std::log(chem->TemperatureStart);
  const double dlnT = (std::log(chem->TemperatureEnd) - lnT0) / (nbin - 1);

  auto fill = [&](const TableSpec &s) {
    double *t = static_cast<double *>(std::malloc(sizeof(double) * nbin));
    if (!t) return false;
    for (int i = 0; i < nbin; ++i) t[i] = s.fn(std::exp(lnT0 + i * dlnT), unit_of[s.unit], cx);
    rates->*(s.slot) = t;
    return true;
  };

  if (chem->primordial_chemistry > 0) {
    for (const TableSpec &s : kPrimordialTables)
      if (!fill(s)) return GR_FAIL;

    // k13dd: 14 fit functions, stored [function][bin]
    rates->k13dd = static_cast<double *>(std::malloc(sizeof(double) * 14 * nbin));
    for (int i = 0; i < nbin; ++i) {
      double f[14];
      r_k13dd(std::exp(lnT0 + i * dlnT), u_k, f);
      for (int j = 0; j < 14; ++j) rates->k13dd[i + nbin * j] = f[j];
    }

    // h2dust: stored [dust bin][gas bin]
    const int nd = chem->NumberOfDustTemperatureBins;
    const double lnTd0 = std::log(chem->DustTemperatureStart);
    const double dlnTd = (std::log(chem->DustTemperatureEnd) - lnTd0) / (nd - 1);
    rates->h2dust = static_cast<double *>(std::malloc(sizeof(double) * nbin * nd));
    for (int i = 0; i < nbin; ++i) {
      const double T = std::exp(lnT0 + i * dlnT);
      for (int j = 0; j < nd; ++j)
        rates->h2dust[i + nbin * j] = r_h2dust(T, std::exp(lnTd0 + j * dlnTd), u_k, cx);
    }
  }

  const bool any_dust = chem->h2_on_dust > 0 || chem->dust_chemistry > 0 ||
                        chem->dust_recombination_cooling > 0;
  if (any_dust)
    for (const TableSpec &s : kDustTables)
      if (!fill(s)) return GR_FAIL;

  rates->comp = 5.65e-36 / u_cool;
  rates->gammah = (chem->photoelectric_heating <= 1) ? chem->photoelectric_heating_rate / u_cool
                                                     : 1.0e-24 / u_cool;
  rates->gamma_isrf = 3.9e-24 / kMh / 0.009387;
  return GR_SUCCESS;
}
