// grackle-b200: per-cell physics as sm_100a device code.  One thread owns one cell and
// keeps its whole state in registers; there are no row temporaries and no masks.
//
// Every function names the reference routine whose result it must reproduce.
// Floating-point operation ORDER is kept where it feeds parity-critical outputs; the
// structure (per-cell, packed table rows, hoisted cell-invariant factors) is ours.
#ifndef GRB200_PHYSICS_CUH
#define GRB200_PHYSICS_CUH

#include "gr_device.h"
#include "fastmath.cuh"

namespace grb {

// FLOAT_8 constants of the Fortran side (phys_const.def, grackle_fortran_types.def)
#define GRB_KBOLTZ 1.3806504e-16
#define GRB_MASS_H 1.67262171e-24
#define GRB_SIGMA_SB 5.670373e-5
#define GRB_TINY 1.0e-20
#define GRB_TINY8 1.0e-40
#define GRB_HUGE8 1.0e40
#define GRB_MU_METAL 16.0
// 1/ln(10) as the reference forms it: 1._DKIND / log(10._DKIND) (cool1d_cloudy_g.F:107)
#define GRB_INV_LN10 (1.0 / 2.302585092994046)

// min / max with the Fortran-like "a < b ? a : b" meaning, spelled in PTX: written in C the
// compiler turns a comparison against a constant into max.f64 / min.f64, whose NaN-aware
// expansion is 9 instructions on sm_100a instead of these 3 (DSETP + 2 FSEL); the species
// floors alone are 12 of them per subcycle.
__device__ __forceinline__ double dmin(double a, double b)
{
  double r;
  asm("{\n\t.reg .pred p;\n\tsetp.lt.f64 p, %1, %2;\n\tselp.f64 %0, %1, %2, p;\n\t}" : "=d"(r) : "d"(a), "d"(b));
  return r;
}
__device__ __forceinline__ double dmax(double a, double b)
{
  double r;
  asm("{\n\t.reg .pred p;\n\tsetp.gt.f64 p, %1, %2;\n\tselp.f64 %0, %1, %2, p;\n\t}" : "=d"(r) : "d"(a), "d"(b));
  return r;
}
__device__ __forceinline__ long long imin(long long a, long long b) { return a < b ? a : b; }
__device__ __forceinline__ long long imax(long long a, long long b) { return a > b ? a : b; }
__device__ __forceinline__ double ld(const double *p) { return __ldg(p); }
__device__ __forceinline__ double sq(double x) { return x * x; }
// x**y for x >= 0 as exp(y ln x): every real power on this path has a non-negative base.
// |y ln x| < 1e2 here, so the result is within ~1e-14 relative of the correctly rounded
// power (the tolerances are 1e-9 / 1e-6) at a fraction of the instructions of pow().
// The elementary functions are the table-driven ones of fastmath.cuh; every caller has the
// kernel parameters `P` in scope, which carry the table pointer.
#define DLOG(x) dlog((x), P.fm_tab)
#define DLOG10(x) dlog10((x), P.fm_tab)
#define DEXP(x) dexp((x), P.fm_tab)
#define DEXP10(x) dexp10((x), P.fm_tab)
#define POW_POS(x, y) dpow_pos((x), (y), P.fm_tab)
__device__ __forceinline__ double pow4(double x) { const double x2 = x * x; return x2 * x2; }

// 1-based clamped index on a uniform axis (interpolators_g.F:28-29 and friends)
__device__ __forceinline__ long long axis_index(double x, const double *par, double dpar, long long n)
{
  return imin(n - 1, imax(1, (long long)((x - ld(par)) / dpar) + 1));
}

// linear interpolation along the last (fastest) axis starting at flat 1-based index `at`
__device__ __forceinline__ double lerp_last(const double *data, long long at, double x,
                                            double x0, double x1)
{
  const double v0 = ld(data + at - 1), v1 = ld(data + at);
  const double slope = (v1 - v0) / (x1 - x0);
  return (x - x0) * slope + v0;
}

// ---- interpolate_1D_g / 2D / 3D / 3Dz / 2Df3D (interpolators_g.F:4-338) ----------
__device__ inline double interp1(const GrCloudyDev &c, const double *data, double in1)
{
  const long long i1 = axis_index(in1, c.par[0], c.dpar[0], c.dim[0]);
  return lerp_last(data, i1, in1, ld(c.par[0] + i1 - 1), ld(c.par[0] + i1));
}

__device__ inline double interp2(const GrCloudyDev &c, const double *data, double in1, double in2)
{
  const long long i1 = axis_index(in1, c.par[0], c.dpar[0], c.dim[0]);
  const long long i2 = axis_index(in2, c.par[1], c.dpar[1], c.dim[1]);
  const double x0 = ld(c.par[1] + i2 - 1), x1 = ld(c.par[1] + i2);
  double v[2];
#pragma unroll
  for (int q = 1; q <= 2; q++)
    v[q - 1] = lerp_last(data, (q + i1 - 2) * c.dim[1] + i2, in2, x0, x1);
  const double slope = (v[1] - v[0]) / (ld(c.par[0] + i1) - ld(c.par[0] + i1 - 1));
  return (in1 - ld(c.par[0] + i1 - 1)) * slope + v[0];
}

__device__ inline double interp3(const GrCloudyDev &c, const double *data, double in1, double in2,
                                 double in3)
{
  const long long i1 = axis_index(in1, c.par[0], c.dpar[0], c.dim[0]);
  const long long i2 = axis_index(in2, c.par[1], c.dpar[1], c.dim[1]);
  const long long i3 = axis_index(in3, c.par[2], c.dpar[2], c.dim[2]);
  const double x0 = ld(c.par[2] + i3 - 1), x1 = ld(c.par[2] + i3);
  const double y0 = ld(c.par[1] + i2 - 1), y1 = ld(c.par[1] + i2);
  double v2[2];
#pragma unroll
  for (int q = 1; q <= 2; q++) {
    double v3[2];
#pragma unroll
    for (int w = 1; w <= 2; w++)
      v3[w - 1] = lerp_last(data, ((q + i1 - 2) * c.dim[1] + (w + i2 - 2)) * c.dim[2] + i3, in3, x0, x1);
    const double slope = (v3[1] - v3[0]) / (y1 - y0);
    v2[q - 1] = (in2 - y0) * slope + v3[0];
  }
  const double slope = (v2[1] - v2[0]) / (ld(c.par[0] + i1) - ld(c.par[0] + i1 - 1));
  return (in1 - ld(c.par[0] + i1 - 1)) * slope + v2[0];
}

// rank dispatch for new-style tables (cool1d_cloudy_g.F:157-236).  Rank-3 (n_H, z, T) grids
// never reach the device: the host collapses them to the (n_H, T) plane at the call's
// redshift (api.cu cloudy_params; interpolate_3Dz_g / interpolate_2Df3D_g semantics), so
// a lookup is at most bilinear.
__device__ inline double cloudy_lookup(const GrCloudyDev &c, const double *data, double log_n_h,
                                       double log10T)
{
  if (c.rank == 1) return interp1(c, data, log10T);
  return interp2(c, data, log_n_h, log10T);
}

// ---- cool1d_cloudy_g (cool1d_cloudy_g.F:4-261): returns the edot increment -------
// Fast form for a rank-2 (n_H, T) table prepared by the host (api.cu cloudy_fast_plane): the
// same bilinear interpolation as interpolate_2D_g (interpolators_g.F:45-101; index clamped,
// weights not, so it extrapolates linearly like the reference), with
//   * cooling and heating read as one (cooling, heating) pair per grid point: 4 gathers serve
//     both lookups instead of 8 gathers + 8 axis gathers,
//   * the weights (x - p_i) / (p_{i+1} - p_i) from one (p_i, 1 / difference) pair per axis, the
//     reciprocal formed on the host.  (Round 1 formed p_i as origin + i * spacing: exact on the
//     synthetic tables, but the real files store rounded axis values - CloudyData_noUVB.h5 has
//     10^1.05 as 11.22018 - and the real-table parity test of round 2 showed a 4e-6 error in the
//     cooling rate from it.)
//   * the CMB-floor term (cooling at T_CMB, cell-invariant along T) read from a per-row array.
// 256-bit read-only load (32-byte aligned address)
__device__ __forceinline__ void ld_nc_v4(const double *p, double &a, double &b, double &c, double &d)
{
  asm("ld.global.nc.v4.f64 {%0, %1, %2, %3}, [%4];" : "=d"(a), "=d"(b), "=d"(c), "=d"(d) : "l"(p));
}

template <class M>
__device__ __forceinline__ double cloudy_edot_fast_m(M &m, const GrParams &P, const GrCloudyDev &c, double rhoH,
                                                     double metallicity, double logtem, int icmbTfloor, int iZscale)
{
  const double log10tem = logtem * GRB_INV_LN10;
  const double log_n_h = m.log10(rhoH * P.dom);
  int i1 = (int)((log_n_h - c.par0[0]) * c.inv_dpar[0]) + 1;
  i1 = min(c.fdim[0] - 1, max(1, i1));
  int i2 = (int)((log10tem - c.par0[1]) * c.inv_dpar[1]) + 1;
  i2 = min(c.fdim[1] - 1, max(1, i2));
  const double2 a1 = __ldg(reinterpret_cast<const double2 *>(c.ax[0]) + (i1 - 1));
  const double2 a2 = __ldg(reinterpret_cast<const double2 *>(c.ax[1]) + (i2 - 1));
  const double w1 = (log_n_h - a1.x) * a1.y;
  const double w2 = (log10tem - a2.x) * a2.y;
#if GRB_CLOUDY_ENTRY == 4
  // one 32-byte entry per grid point: (cooling, heating) of the point AND of its right-hand T
  // neighbour, so the four corners are two 256-bit loads (sm_100a LDG.E.256) instead of four 128-bit
  // ones.  Every lane reads its own (n_H, T) point - 32 different lines per instruction - and both
  // kernels that call this are bound by the wavefronts of the load/store data pipe (ncu round 2:
  // 86 % / 90 % of peak; these gathers were 25 tag requests each, a quarter of the kernel's total).
  const double *g = c.ch + 4 * ((i1 - 1) * c.fdim[1] + (i2 - 1));
  double2 v00, v01, v10, v11;
  ld_nc_v4(g, v00.x, v00.y, v01.x, v01.y);
  ld_nc_v4(g + 4 * c.fdim[1], v10.x, v10.y, v11.x, v11.y);
#else
  const double2 *g = reinterpret_cast<const double2 *>(c.ch) + ((i1 - 1) * c.fdim[1] + (i2 - 1));
  const double2 v00 = __ldg(g), v01 = __ldg(g + 1), v10 = __ldg(g + c.fdim[1]), v11 = __ldg(g + c.fdim[1] + 1);
#endif
  const bool floor_on = (icmbTfloor == 1) && ((log10tem - P.log10_tcmb) < 2.0);
  double m0 = 0.0, m1 = 0.0;
  if (floor_on) { m0 = ld(c.cmb + i1 - 1); m1 = ld(c.cmb + i1); }
  const double c0 = w2 * (v01.x - v00.x) + v00.x, c1 = w2 * (v11.x - v10.x) + v10.x;
  // cooling and heating: two independent powers of ten, formed side by side; the CMB-floor term
  // only where the gas is within two decades of T_CMB (few cells: a branch, not a third power)
  const double e_cool = m.exp10(w1 * (c1 - c0) + c0);
  const double h0 = w2 * (v01.y - v00.y) + v00.y, h1 = w2 * (v11.y - v10.y) + v10.y;
  const double e_heat = (c.get_heat == 1) ? m.exp10(w1 * (h1 - h0) + h0) : 0.0;
  double edot_met = -e_cool;
  if (floor_on) edot_met = edot_met + m.exp10(w1 * (m1 - m0) + m0);
  if (c.get_heat == 1) edot_met = edot_met + e_heat;
  if (iZscale == 1) edot_met = edot_met * metallicity;
  return edot_met * rhoH * rhoH;
}

__device__ __forceinline__ double cloudy_edot_fast(const GrParams &P, const GrCloudyDev &c, double rhoH,
                                                   double metallicity, double logtem, int icmbTfloor,
                                                   int iZscale)
{
  FmFast f{P.fm_tab, false};
  double v = cloudy_edot_fast_m(f, P, c, rhoH, metallicity, logtem, icmbTfloor, iZscale);
  if (f.bad) {       // an argument outside the lean functions' range: once more with the checked forms
    FmChecked e{P.fm_tab};
    v = cloudy_edot_fast_m(e, P, c, rhoH, metallicity, logtem, icmbTfloor, iZscale);
  }
  return v;
}

__device__ inline double cloudy_edot(const GrParams &P, const GrCloudyDev &c, double rhoH,
                                     double metallicity, double logtem, int icmbTfloor, int iZscale)
{
  if (c.fast == 1) return cloudy_edot_fast(P, c, rhoH, metallicity, logtem, icmbTfloor, iZscale);
  const double log10_tCMB = P.log10_tcmb;
  const double log10tem = logtem * GRB_INV_LN10;
  const double log_n_h = DLOG10(rhoH * P.dom);
  double edot_met = -DEXP10(cloudy_lookup(c, c.cooling, log_n_h, log10tem));
  if (icmbTfloor == 1 && (log10tem - log10_tCMB) < 2.0)
    edot_met = edot_met + DEXP10(cloudy_lookup(c, c.cooling, log_n_h, log10_tCMB));
  if (c.get_heat == 1)
    edot_met = edot_met + DEXP10(cloudy_lookup(c, c.heating, log_n_h, log10tem));
  if (iZscale == 1) edot_met = edot_met * metallicity;
  return edot_met * rhoH * rhoH;
}

// ---- interpolate_4D_g / interpolate_5D_g (interpolators_g.F:343-566) ----------------
// One routine for any rank R: axis q is interpolated from the two results of axis q+1, the
// last (fastest) axis from two neighbouring data values - the nesting, index clamping
// (uniform-axis index clamped, weight not) and operation order of the reference's unrolled
// rank-4 and rank-5 routines.  `flat` is the 0-based flat index of the corner chosen so far.
// The axes travel BY VALUE in a small struct and the routine is a real (non-inlined) call:
//  * a reference into the kernel's by-value parameter copy would take the address of that
//    copy and send ALL of it - in every instantiation of the solve kernel, also where this
//    code is dead - to local memory (measured: 1952-byte stack frame, compact kernel 2.15x
//    slower);
//  * inlined, the 31 nested interpolations of three lookups cost the general instantiation
//    up to 19 % (C2) through code size and register pressure, for a table format almost no
//    run uses.
struct InterpAxes {
  const double *par[5];
  long long dim[5];
  double dpar[5];
};

template <int R, int Q>
__device__ __forceinline__ double interp_axis(const InterpAxes &c, const double *data, const double (&in)[R],
                                              const long long (&ix)[R], long long flat)
{
  const double x0 = ld(c.par[Q] + ix[Q] - 1), x1 = ld(c.par[Q] + ix[Q]);
  double v0, v1;
  if constexpr (Q == R - 1) {
    const long long at = flat * c.dim[Q] + ix[Q];     // 1-based position on the last axis
    v0 = ld(data + at - 1);
    v1 = ld(data + at);
  } else {
    v0 = interp_axis<R, Q + 1>(c, data, in, ix, flat * c.dim[Q] + ix[Q] - 1);
    v1 = interp_axis<R, Q + 1>(c, data, in, ix, flat * c.dim[Q] + ix[Q]);
  }
  const double slope = (v1 - v0) / (x1 - x0);
  return (in[Q] - x0) * slope + v0;
}

template <int R>
__device__ __noinline__ double interp_rank(const InterpAxes c, const double *data, double in0, double in1,
                                           double in2, double in3, double in4)
{
  const double all[5] = {in0, in1, in2, in3, in4};
  double in[R];
  long long ix[R];
#pragma unroll
  for (int q = 0; q < R; q++) {
    in[q] = all[q];
    ix[q] = axis_index(in[q], c.par[q], c.dpar[q], c.dim[q]);
  }
  return interp_axis<R, 0>(c, data, in, ix, 0);
}

__device__ __forceinline__ InterpAxes interp_axes_of(const GrCloudyDev &c)
{
  InterpAxes a;
#pragma unroll
  for (int q = 0; q < 5; q++) { a.par[q] = c.par[q]; a.dim[q] = c.dim[q]; a.dpar[q] = c.dpar[q]; }
  return a;
}

// ---- old-style tables, any rank 1..5 (cool1d_cloudy_old_tables_g.F:4-335).  Rank 4 adds the
// electron-fraction axis, rank 5 the redshift axis in front of the temperature
// (:138-146, :246-318); `de` is only read for those ranks. ---------------------------
__device__ inline double cloudy_old_edot(const GrParams &P, const GrCloudyDev &c, double d, double de,
                                         double rhoH, double metallicity, double logtem)
{
  const double log10_tCMB = P.log10_tcmb;        // see oracle note on the unassigned upstream local
  const double log10tem = logtem * GRB_INV_LN10;
  double log_n_h = 0.0, log_Z = 0.0, log_e_frac = 0.0, cl_e_frac = 1.0;
  if (c.rank > 1) log_n_h = DLOG10(rhoH * P.dom);
  if (c.rank > 2) log_Z = DLOG10(metallicity);
  const bool rank45 = (P.iold45 != 0) && c.rank > 3;   // compiled only into the VAR_RANK45 instantiations
  if (rank45) {                                  // :138-146
    const double fh = rhoH / d;
    const double e_frac = 2.0 * de / (d * (1.0 + fh));
    log_e_frac = dmin(log10(e_frac), 0.0);       // e_frac = 0 must give -Inf: library log10
    cl_e_frac = e_frac * (1.0 + (2.0 * P.clEleFra * metallicity * fh) / (1.0 + fh));
  }
  auto look = [&](const double *data, double lt) {
    if (c.rank == 1) return interp1(c, data, lt);
    if (c.rank == 2) return interp2(c, data, log_n_h, lt);
    if (c.rank == 3 || !rank45) return interp3(c, data, log_n_h, log_Z, lt);
    if (c.rank == 4) return interp_rank<4>(interp_axes_of(c), data, log_n_h, log_Z, log_e_frac, lt, 0.0);
    return interp_rank<5>(interp_axes_of(c), data, log_n_h, log_Z, log_e_frac, P.zr, lt);
  };
  double edot_met = -DEXP10(look(c.cooling, log10tem));
  if (P.icmbTfloor == 1 && (log10tem - log10_tCMB) < 2.0)
    edot_met = edot_met + DEXP10(look(c.cooling, log10_tCMB));
  if (P.iClHeat == 1) edot_met = edot_met + DEXP10(look(c.heating, log10tem));
  if (rank45) edot_met = edot_met * cl_e_frac;
  return edot_met * rhoH * d;
}

// ---- calc_temp1d_cloudy_g (calc_temp1d_cloudy_g.F:150-242): T and mu for pc=0 -----
__device__ inline void temperature_table(const GrParams &P, double d, double e, double metal,
                                         double rhoH, double &tgas, double &mmw)
{
  const GrCloudyDev &c = P.pri;
  const double inv_log10 = GRB_INV_LN10;
  const double log_n_h = DLOG10(rhoH * P.dom);
  double munew = 1.0, muold = 1.0;
  for (int ti = 1; ti <= 20; ti++) {
    muold = munew;
    tgas = dmax((P.gamma - 1.0) * e * munew * P.utem, P.temstart);
    const double log10tem = DLOG(tgas) * inv_log10;
    munew = cloudy_lookup(c, c.mmw, log_n_h, log10tem);
    munew = 0.5 * (munew + muold);
    tgas = tgas * munew / muold;
    if (fabs((munew / muold) - 1.0) <= 1.e-2) {
      muold = munew;
      if (P.imetal == 1) {
        munew = d / ((d - metal) / munew + metal / GRB_MU_METAL);
        tgas = tgas * munew / muold;
      }
      break;
    }
  }
  mmw = munew;
}

// ---- calc_tdust_1d_g + calc_kappa_gr_g + calc_gr_balance_g (calc_tdust_1d_g.F) ----
__device__ __forceinline__ double kappa_gr(double td)
{
  if (td < 200.0) return 4.0e-4 * (td * td);
  if (td < 1.5e3) return 16.0;
  const double x = ddiv(td, 1.5e3);
  const double x3 = (x * x) * x;
  const double x6 = x3 * x3;
  return dmax(GRB_TINY, 16.0 * ddiv_s(1.0, x6 * x6));
}
__device__ __forceinline__ double gr_balance(double td, double tgas, double kgr, double trad4,
                                             double gasgr, double gamma_isrf, double nh)
{
  return gamma_isrf + (4.0 * GRB_SIGMA_SB) * kgr * (trad4 - pow4(td)) + (gasgr * nh * (tgas - td));
}

__device__ inline double dust_temperature(const GrParams &P, double tgas, double nh, double gasgr,
                                          double gamma_isrfa, double isrf, double trad_in)
{
  const double t_subl = 1.5e3, radf = 4.0 * GRB_SIGMA_SB, kgr1 = 4.0e-4;
  const double trad = dmax(1.0, trad_in);
  const double trad4 = pow4(trad);
  const double gamma_isrf = isrf * gamma_isrfa;
  bool newton = true, bisect = true;
  double td = 0.0, pert = 1.e-3;
  if (trad >= tgas) {
    td = trad; newton = false; bisect = false;
  } else if (tgas > t_subl) {
    newton = false;
  } else {
    td = dmax(trad, POW_POS(gamma_isrf / radf / kgr1, 0.17));
  }
  for (int it = 1; it <= 50 && newton; it++) {
    const double tdp = dmax(1.e-3, ((1.0 + pert) * td));
    const double sol = gr_balance(td, tgas, kappa_gr(td), trad4, gasgr, gamma_isrf, nh);
    const double solp = gr_balance(tdp, tgas, kappa_gr(tdp), trad4, gasgr, gamma_isrf, nh);
    // checked lean divisions (<= 1 ulp; a vanishing or non-finite denominator takes the IEEE
    // division, so a flat slope still sends td to +-Inf and hands over to the bisection): the
    // plain '/' here was 20 % of the instructions of config C3's subcycle
    const double slope = ddiv_s(solp - sol, pert * td);
    const double told = td;
    td = td - ddiv_s(sol, slope);
    pert = dmax(dmin(pert, ddiv_s(0.5 * fabs(td - told), td)), 1.e-10);
    if (td < trad) newton = false;                       // give up, bisection takes over
    // |sol / solp| < 1e-5 as a product: the quotient is only compared, never used (the two forms can differ
    // only when the quotient is within an ulp of the threshold)
    else if (fabs(sol) < 1.e-5 * fabs(solp)) { newton = false; bisect = false; }
  }
  if (bisect) {
    td = trad;
    double hi = tgas;
    for (int it = 1; it <= 30; it++) {
      double mid = 0.5 * (td + hi);
      if (it == 1) mid = dmin(mid, t_subl);
      const double sol = gr_balance(mid, tgas, kappa_gr(mid), trad4, gasgr, gamma_isrf, nh);
      if (sol > 0.0) td = mid; else hi = mid;
      if (fabs(hi - td) <= 1.e-3 * td) break;          // (td >= trad >= 1: the division of the reference as a product)
    }
  }
  return td;
}

}  // namespace grb
#endif
