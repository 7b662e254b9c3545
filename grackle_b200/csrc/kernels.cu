// grackle-b200: the sm_100a kernels.  One thread per active cell, whole cell state in
// registers, SoA fields read/written exactly once per call, comoving<->proper scaling
// and species floors/renormalisation fused into the same pass.
//
//   k_pressure / k_temperature / k_gamma   calculate_{pressure,temperature,gamma}.c   HBM-bound
//   k_temperature_table                    calc_temp_cloudy_g.F (pc = 0)              HBM + L2 gathers
//   k_cooling_time                         cool_multi_time_g.F                        FP64 + HBM
//   k_dust_temperature                     calc_tdust_3d_g.F                          FP64 (Newton)
//   k_solve                                solve_rate_cool_g.F (whole file)           FP64 pipe / instruction issue
//
// No tensor cores: there is no dense contraction anywhere on this path (north_star).
#include <cuda_runtime.h>

#include "network.cuh"
#include "gr_launch.h"

namespace grb {

// active-cell id -> flat field index (i fastest), plus (i,j,k) when needed
__device__ __forceinline__ size_t cell_index(const GrParams &P, long long t, int &i, int &j, int &k)
{
  const int ni = P.ie - P.is + 1, nj = P.je - P.js + 1;
  if (P.nactive <= 0x7fffffffLL) {      // 32-bit divisions (1024^3 = 2^30 cells still qualifies)
    const unsigned tt = (unsigned)t, r = tt / (unsigned)ni;
    i = (int)(tt - r * (unsigned)ni) + P.is;
    const unsigned q = r / (unsigned)nj;
    j = (int)(r - q * (unsigned)nj) + P.js;
    k = (int)q + P.ks;
  } else {
    i = (int)(t % ni) + P.is;
    const long long r = t / ni;
    j = (int)(r % nj) + P.js;
    k = (int)(r / nj) + P.ks;
  }
  return (size_t)i + (size_t)P.in * ((size_t)j + (size_t)P.jn * (size_t)k);
}

// Load one cell, converting comoving -> proper in registers (scale_fields_g) and applying
// the species floors (ceiling_species_g) when `floors` is set (solve only).
template <int PC, bool SCALE_DIV>
__device__ __forceinline__ void load_cell(const GrParams &P, size_t x, Cell &c, bool floors)
{
  const double s = (P.iexpand == 1) ? P.a3inv : 1.0;
  const bool sc = (P.iexpand == 1);
  // solve multiplies by aye**(-3) (solve_rate_cool_g.F:353); cooling time divides by aye**3
  // (cool_multi_time_g.F:163): keep each spelling so roundings match
#define LDF(id) (sc ? (SCALE_DIV ? ddiv(P.f[id][x], P.a3) : P.f[id][x] * s) : P.f[id][x])
  c.d = LDF(F_D);
  c.e = P.f[F_E][x];
  c.metal = (P.imetal == 1) ? LDF(F_METAL) : 0.0;
  c.dust = (P.idustfield == 1) ? LDF(F_DUST) : 0.0;
  if (PC > 0) {
    c.de = LDF(F_DE); c.HI = LDF(F_HI); c.HII = LDF(F_HII);
    c.HeI = LDF(F_HEI); c.HeII = LDF(F_HEII); c.HeIII = LDF(F_HEIII);
  } else {
    c.de = c.HI = c.HII = c.HeI = c.HeII = c.HeIII = 0.0;
  }
  if (PC > 1) { c.HM = LDF(F_HM); c.H2I = LDF(F_H2I); c.H2II = LDF(F_H2II); }
  else { c.HM = c.H2I = c.H2II = 0.0; }
  if (PC > 2) { c.DI = LDF(F_DI); c.DII = LDF(F_DII); c.HDI = LDF(F_HDI); }
  else { c.DI = c.DII = c.HDI = 0.0; }
#undef LDF
  if (floors) {
    if (PC > 0) {
      c.de = dmax(c.de, GRB_TINY); c.HI = dmax(c.HI, GRB_TINY); c.HII = dmax(c.HII, GRB_TINY);
      c.HeI = dmax(c.HeI, GRB_TINY); c.HeII = dmax(c.HeII, GRB_TINY);
      c.HeIII = dmax(c.HeIII, 1e-5 * GRB_TINY);
    }
    if (PC > 1) { c.HM = dmax(c.HM, GRB_TINY); c.H2I = dmax(c.H2I, GRB_TINY); c.H2II = dmax(c.H2II, GRB_TINY); }
    if (PC > 2) { c.DI = dmax(c.DI, GRB_TINY); c.DII = dmax(c.DII, GRB_TINY); c.HDI = dmax(c.HDI, GRB_TINY); }
    if (P.imetal == 1) c.metal = dmax(c.metal, GRB_TINY);
  }
  c.vheat = (P.iVheat == 1) ? P.f[F_VHEAT][x] : 0.0;
  c.mheat = (P.iMheat == 1) ? P.f[F_MHEAT][x] : 0.0;
  c.isrf = (P.iisrffield > 0) ? P.f[F_ISRF][x] : 0.0;
  if (P.iradtrans == 1) {
    c.kphHI = P.f[F_KPHHI][x]; c.kphHeI = P.f[F_KPHHEI][x]; c.kphHeII = P.f[F_KPHHEII][x];
    c.photogamma = P.f[F_PHOTOGAMMA][x];
    c.kdissH2I = (PC > 1) ? P.f[F_KDISSH2I][x] : 0.0;
  } else {
    c.kphHI = c.kphHeI = c.kphHeII = c.photogamma = c.kdissH2I = 0.0;
  }
  c.xH2shield = (PC > 1 && P.iH2shield == 2) ? P.f[F_XH2SHIELD][x] : 0.0;
  c.fshield = (PC > 1 && P.iH2shieldcustom > 0) ? P.f[F_FSHIELD_CUSTOM][x] : 1.0;
}

// ------------------------------------------------------------------------------
// calculate_pressure / temperature / gamma  (pc > 0; proper or comoving alike: these
// routines never rescale, calculate_pressure.c:31-128, calculate_temperature.c:64-148,
// calculate_gamma.c:37-124).  MODE 0 = pressure, 1 = temperature, 2 = gamma.
// ------------------------------------------------------------------------------
template <int PC, int MODE>
__global__ void __launch_bounds__(256) k_thermo(const __grid_constant__ GrParams P)
{
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= P.nactive) return;
  int i, j, k;
  const size_t x = cell_index(P, t, i, j, k);
  const double tiny_number = 1.e-20;
  const double d = P.f[F_D][x], e = P.f[F_E][x];
  double p = ((P.gamma - 1.0) * d * e);
  if (p < tiny_number) p = tiny_number;

  double HI = 0, HII = 0, HeI = 0, HeII = 0, HeIII = 0, de = 0, HM = 0, H2I = 0, H2II = 0;
  if (PC > 0 && (MODE > 0 || PC > 1)) {
    HI = P.f[F_HI][x]; HII = P.f[F_HII][x]; HeI = P.f[F_HEI][x]; HeII = P.f[F_HEII][x];
    HeIII = P.f[F_HEIII][x]; de = P.f[F_DE][x];
  }
  if (PC > 1) { HM = P.f[F_HM][x]; H2I = P.f[F_H2I][x]; H2II = P.f[F_H2II][x]; }

  if (PC > 1) {   // H2 vibrational correction of the pressure
    const double GammaInverse = P.inv_gm1;
    double number_density = 0.25 * (HeI + HeII + HeIII) + HI + HII + HM + de;
    const double nH2 = 0.5 * (H2I + H2II);
    if (number_density == 0) number_density = tiny_number;
    const double temp = dmax(ddiv_s(P.utem * p, number_density + nH2), 1);
    double GammaH2Inverse = 0.5 * 5.0;
    if (ddiv_s(nH2, number_density) > 1e-3) {
      const double xx = ddiv_s(6100.0, temp);
      if (xx < 10.0) {
        const double ex = exp(xx);
        GammaH2Inverse = 0.5 * (5 + 2.0 * xx * xx * ex / ((ex - 1.0) * (ex - 1.0)));
      }
    }
    const double Gamma1 = 1.0 + ddiv_s(nH2 + number_density, nH2 * GammaH2Inverse + number_density * GammaInverse);
    p *= (Gamma1 - 1.0) * P.inv_gm1;
  }
  if (MODE == 0) { P.out[x] = p; return; }

  // temperature (quirk Q1: the upstream density floor is the constant 1.-20 = -19, i.e. inert)
  double number_density = 0.25 * (HeI + HeII + HeIII) + HI + HII + de;
  if (PC > 1) number_density += HM + 0.5 * (H2I + H2II);
  if (P.imetal == 1) number_density += P.f[F_METAL][x] * (1.0 / GRB_MU_METAL);
  double T = p * ddiv_s(P.utem, dmax(number_density, 1. - 20));
  T = dmax(T, 1.0);
  if (MODE == 1) { P.out[x] = T; return; }

  // gamma (only reached for PC > 1; PC <= 1 is a host-side fill)
  const double GammaInverse = P.inv_gm1;
  number_density = 0.25 * (HeI + HeII + HeIII) + HI + HII + HM + de;
  const double nH2 = 0.5 * (H2I + H2II);
  double GammaH2Inverse = 0.5 * 5.0;
  if (ddiv_s(nH2, number_density) > 1e-3) {
    const double xx = ddiv_s(6100.0, T);
    if (xx < 10.0) {
      const double ex = exp(xx);
      GammaH2Inverse = 0.5 * (5 + 2.0 * xx * xx * ex / ((ex - 1.0) * (ex - 1.0)));
    }
  }
  P.out[x] = 1.0 + ddiv_s(nH2 + number_density, nH2 * GammaH2Inverse + number_density * GammaInverse);
}

// constant fill of the active region (calculate_gamma for pc <= 1)
__global__ void __launch_bounds__(256) k_fill(const __grid_constant__ GrParams P, double value)
{
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= P.nactive) return;
  int i, j, k;
  P.out[cell_index(P, t, i, j, k)] = value;
}

// calc_temp_cloudy_g.F:141-171 (pc = 0).  NB rhoH = fh*d here, without the metal
// subtraction cool1d_multi_g applies.
__global__ void __launch_bounds__(256) k_temperature_table(const __grid_constant__ GrParams P)
{
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= P.nactive) return;
  int i, j, k;
  const size_t x = cell_index(P, t, i, j, k);
  const double s = (P.iexpand == 1) ? P.a3inv : 1.0;
  const double d = (P.iexpand == 1) ? P.f[F_D][x] * s : P.f[F_D][x];
  const double metal = (P.imetal == 1) ? ((P.iexpand == 1) ? P.f[F_METAL][x] * s : P.f[F_METAL][x]) : 0.0;
  double tgas, mmw;
  temperature_table(P, d, P.f[F_E][x], metal, P.fh * d, tgas, mmw);
  P.out[x] = tgas;
}

// cool_multi_time_g.F: one cool_rate evaluation, t_cool = max(p/(gamma-1), tiny)/edot
template <int VAR>
__host__ __device__ inline void variant_flags(GrParams &P);

// calc_tdust_3d_g.F: gas temperature comes in through P.f[F_VHEAT] (scratch slot)
template <int PC>
__global__ void __launch_bounds__(128) k_dust_temperature(const __grid_constant__ GrParams P)
{
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= P.nactive) return;
  int i, j, k;
  const size_t x = cell_index(P, t, i, j, k);
  const double myisrf = (P.iisrffield > 0) ? P.f[F_ISRF][x] : P.isrf;
  double nh = P.f[F_HI][x] + P.f[F_HII][x];
  if (PC > 1) nh = nh + P.f[F_H2I][x] + P.f[F_H2II][x];
  nh = nh * P.urho / GRB_MASS_H;
  const double tgas = P.f[F_VHEAT][x];
  double logtem = log(tgas);
  logtem = dmax(logtem, P.logtem0);
  logtem = dmin(logtem, P.logtem9);
  const TIndex ti = t_index(P, logtem);
  const double *row = P.cool + (size_t)(ti.idx - 1) * (NCOOL * 2);
  double gasgr = lerp_row(row, L_GASGR, ti.tdef);
  gasgr = gasgr * P.fgr * P.coolunit / GRB_MASS_H;
  P.out[x] = dust_temperature(P, tgas, nh, gasgr, P.gamma_isrfa, myisrf, P.comp2);
}

// ------------------------------------------------------------------------------
// k_solve: solve_rate_cool_g.F for every active cell.
//
// Divergence control ("lane refill"): subcycle counts differ by orders of magnitude
// between cells, so a static thread<->cell mapping would leave most lanes of a warp idle
// while the slowest cell finishes (the reference's OpenMP rows have the same problem).
// Here a lane that finishes its cell immediately takes the next unprocessed cell from a
// global queue (one warp-aggregated atomicAdd per refill, found with __ballot_sync), so
// every lane executes useful subcycles until the queue is empty.  Each cell keeps its own
// iteration counter, hence the iter==1 / iter>50 heuristics of the reference are
// reproduced exactly and results do not depend on which lane processed a cell.
// ------------------------------------------------------------------------------
struct SubState {           // per-cell carry between subcycles
  double ttot, tgasold, dedot_prev, HIdot_prev;
  int iter;
};

// One subcycle of solve_rate_cool_g.F:451-817 for one cell is split into per-cell stages
// with the warp-cooperative table fetches between them (see k_solve): thermo_head, [fetch
// A], cool_sums + k13_martin, [fetch B], subcycle_lookup, subcycle_finish (rate-limited dt,
// energy update, species sweep).  subcycle_finish returns true when the cell has reached dt.
template <int PC>
__device__ __forceinline__ void subcycle_lookup(const GrParams &P, const Cell &c, const Thermo &th, size_t x,
                                                const TIndex &ti, double k13dd, const Rates &r)
{
  Neigh nb;
  if (PC > 1 && P.iH2shield == 1) {
    const double sc = (P.iexpand == 1) ? P.a3inv : 1.0;
    const size_t sj = (size_t)P.in, sk = (size_t)P.in * P.jn;
    const double *D = P.f[F_D];
    nb.d[0] = D[x + 1] * sc; nb.d[1] = D[x - 1] * sc; nb.d[2] = D[x + sj] * sc;
    nb.d[3] = D[x - sj] * sc; nb.d[4] = D[x + sk] * sc; nb.d[5] = D[x - sk] * sc;
  }
  lookup_rates<PC>(P, c, th, nb, ti, k13dd, r);
}

template <int PC>
__device__ __forceinline__ bool subcycle_finish(const GrParams &P, Cell &c, SubState &s, const Thermo &th,
                                                const TIndex &ti, const Rates &r)
{
  const double dt = P.dt;
  double dtit = GRB_HUGE8;
  double edot = th.edot;

  if (PC > 0) {
    double dedot, HIdot;
    rate_derivs<PC>(P, c, th, r, dedot, HIdot, edot);

    // solve_rate_cool_g.F:554-646
    if (fabs(dedot) < GRB_TINY8) dedot = dmin(GRB_TINY, c.de);
    if (fabs(HIdot) < GRB_TINY8) HIdot = dmin(GRB_TINY, c.HI);
    if (dmin(fabs(RK(k1) * c.de * c.HI), fabs(RK(k2) * c.HII * c.de)) > 1.0e6 * dmax(fabs(dedot), fabs(HIdot))) {
      dedot = GRB_TINY8;
      HIdot = GRB_TINY8;
    }
    if (s.iter > 50) {
      dedot = dmin(fabs(dedot), fabs(s.dedot_prev));
      HIdot = dmin(fabs(HIdot), fabs(s.HIdot_prev));
    }
    // (after iter > 50 either rate can be exactly zero: the quotient must be +Inf then)
    dtit = dmin(dmin(fabs(ddiv_s(0.1 * c.de, dedot)), fabs(ddiv_s(0.1 * c.HI, HIdot))), dmin(dt - s.ttot, 0.5 * dt));

    if (PC > 1 && c.d * P.dom > 1e8 && edot > 0.0) {
      // keep the 3-body equilibrium H abundance from changing by more than 10 %
      const double fd = P.fh * c.d;
      const int ridx = ti.idx;
      const double lt1 = (P.logtem0 + (ridx - 1) * P.dlogtem), lt2 = (P.logtem0 + (ridx) * P.dlogtem);   // = t_index
      const double k22lo = ld(P.k22a + ridx - 1), k22hi = ld(P.k22a + ridx);
      const double k13lo = ld(P.k13a + ridx - 1), k13hi = ld(P.k13a + ridx);
      const double eqt2 = dmin(th.logtgas + 0.1 * P.dlogtem, lt2);
      double eqtdef = (eqt2 - lt1) * P.inv_dlogtem;
      const double eqk222 = k22lo + (k22hi - k22lo) * eqtdef;
      const double eqk132 = k13lo + (k13hi - k13lo) * eqtdef;
      const double heq2 = ddiv(-1.0, 4.0 * eqk222) * (eqk132 - sqrt(8.0 * eqk132 * eqk222 * fd + eqk132 * eqk132));
      const double eqt1 = dmax(th.logtgas - 0.1 * P.dlogtem, lt1);
      eqtdef = (eqt1 - lt1) * P.inv_dlogtem;
      const double eqk221 = k22lo + (k22hi - k22lo) * eqtdef;
      const double eqk131 = k13lo + (k13hi - k13lo) * eqtdef;
      const double heq1 = ddiv(-1.0, 4.0 * eqk221) * (eqk131 - sqrt(8.0 * eqk131 * eqk221 * fd + eqk131 * eqk131));
      const double dheq = ddiv(fabs(heq2 - heq1), DEXP(eqt2) - DEXP(eqt1)) * ddiv(th.tgas, th.p2d) * edot;
      const double heq = ddiv(-1.0, 4.0 * RK(k22)) * (RK(k13) - sqrt(8.0 * RK(k13) * RK(k22) * fd + RK(k13) * RK(k13)));
      dtit = dmin(dtit, ddiv(0.1 * heq, dheq));
    }
    // (the reference's iter>10 rule min(1.5*olddtit, dtit) is inert: olddtit is always huge8)
  }

  // cooling/heating limit (:698-750)
  const double energy = dmax(th.p2d * P.inv_gm1, GRB_TINY8);
  if (th.tgas <= 1.01 * P.temstart && edot < 0.0) edot = GRB_TINY8;
  if (fabs(edot) < GRB_TINY8) edot = GRB_TINY8;
  dtit = dmin(dmin(fabs(ddiv(0.1 * energy, edot)), dt - s.ttot), dtit);

  if (P.icool == 1) c.e = c.e + ddiv(edot, c.d) * dtit;
  if (PC > 0) step_species<PC>(P, c, th, r, dtit, s.HIdot_prev, s.dedot_prev);

  s.ttot = dmin(s.ttot + dtit, dt);
  return fabs(dt - s.ttot) < 1.0e-10 * dt;
}

// extras that are only read when their physics is on are re-fetched per subcycle so they
// do not occupy registers across the loop
template <int PC>
__device__ __forceinline__ void load_extras(const GrParams &P, size_t x, Cell &c)
{
  c.vheat = (P.iVheat == 1) ? P.f[F_VHEAT][x] : 0.0;
  c.mheat = (P.iMheat == 1) ? P.f[F_MHEAT][x] : 0.0;
  c.isrf = (P.iisrffield > 0) ? P.f[F_ISRF][x] : 0.0;
  if (P.iradtrans == 1) {
    c.kphHI = P.f[F_KPHHI][x]; c.kphHeI = P.f[F_KPHHEI][x]; c.kphHeII = P.f[F_KPHHEII][x];
    c.photogamma = P.f[F_PHOTOGAMMA][x];
    c.kdissH2I = (PC > 1) ? P.f[F_KDISSH2I][x] : 0.0;
  } else {
    c.kphHI = c.kphHeI = c.kphHeII = c.photogamma = c.kdissH2I = 0.0;
  }
  c.xH2shield = (PC > 1 && P.iH2shield == 2) ? P.f[F_XH2SHIELD][x] : 0.0;
  c.fshield = (PC > 1 && P.iH2shieldcustom > 0) ? P.f[F_FSHIELD_CUSTOM][x] : 1.0;
}

// finished cell -> memory: proper -> comoving, make_consistent_g, store species + energy
template <int PC>
__device__ __forceinline__ void store_cell(const GrParams &P, size_t x, Cell &c)
{
  if (P.iexpand == 1) {
    const double s = P.a3;
    c.d *= s; c.metal *= s;
    if (PC > 0) { c.de *= s; c.HI *= s; c.HII *= s; c.HeI *= s; c.HeII *= s; c.HeIII *= s; }
    if (PC > 1) { c.HM *= s; c.H2I *= s; c.H2II *= s; }
    if (PC > 2) { c.DI *= s; c.DII *= s; c.HDI *= s; }
  }
  if (PC > 0) make_consistent<PC>(P, c);
  P.f[F_E][x] = c.e;
  if (PC > 0) {
    P.f[F_DE][x] = c.de; P.f[F_HI][x] = c.HI; P.f[F_HII][x] = c.HII;
    P.f[F_HEI][x] = c.HeI; P.f[F_HEII][x] = c.HeII; P.f[F_HEIII][x] = c.HeIII;
  }
  if (PC > 1) { P.f[F_HM][x] = c.HM; P.f[F_H2I][x] = c.H2I; P.f[F_H2II][x] = c.H2II; }
  if (PC > 2) { P.f[F_DI][x] = c.DI; P.f[F_DII][x] = c.DII; P.f[F_HDI][x] = c.HDI; }
}

// CTA-wide re-alignment points.  The executed path of one subcycle (31 KB of SASS in the
// compact instantiation, 66 KB before round 1's code-size work) competes with a 32 KB
// instruction cache; warps that drift apart each stream it from L2 on their own (first
// version: 55 % of all stall cycles were stall_no_inst).  A barrier per subcycle keeps the
// warps of a group inside the same few KB of code, so one warp's fetch serves all of them.
// The CTA is split into lockstep groups of GRB_GROUP_WARPS warps, each on its own named
// barrier.  With groups of 4 every SM sub-partition holds one warp of each group: the four
// warps a scheduler sees are in different phases and cover each other's latencies.
// Measured (Mcell/s, C4 sample): 1 warp 31.2, 2: 32.3, 4: 31.9, 8: 29.0, 16: 21.8.
// The best group size depends on how much code a subcycle streams: the compact instantiation
// (5.6k instructions) is fastest with 2 warps per group (C1 +3 %, C2 +10 %, C4 +2 % over 4),
// the larger dust + shielding and general ones with 4 (C3 -5 % with 2).
#ifndef GRB_GROUP_WARPS
#define GRB_GROUP_WARPS 4
#endif
#ifndef GRB_GROUP_WARPS_LEAN
#define GRB_GROUP_WARPS_LEAN 2
#endif
// ... and the compact kernels of the two smallest networks (pc = 0, 1) with no grouping at all
// (C1 +3 %, C2 +9 % over groups of 2; C4 -4 %)
#ifndef GRB_GROUP_WARPS_LEAN_SMALL
#define GRB_GROUP_WARPS_LEAN_SMALL 1
#endif
// ... and the general / dust + shielding kernels of the H2 networks (pc = 2, 3: the most code per
// subcycle) with 8 (general C3 +9 %, C4 +4 %, dust + shielding C3 +1.5 % over 4; pc = 0/1 -2 / -10 %)
#ifndef GRB_GROUP_WARPS_BIG
#define GRB_GROUP_WARPS_BIG 8
#endif
static_assert(GRB_SOLVE_THREADS % (32 * GRB_GROUP_WARPS_BIG) == 0, "CTA must be whole lockstep groups");
// lanes per cell in the cooperative table fetches (fewer lanes = more loads per lane and pass)
#ifndef GRB_G_MAIN
#define GRB_G_MAIN 8
#endif
#ifndef GRB_G_COOL
#define GRB_G_COOL 8
#endif
#ifndef GRB_G_DD
#define GRB_G_DD 4
#endif

static_assert(GRB_SOLVE_THREADS % (32 * GRB_GROUP_WARPS) == 0 && GRB_SOLVE_THREADS % (32 * GRB_GROUP_WARPS_LEAN) == 0,
              "CTA must be whole lockstep groups");
static_assert((GRB_GROUP_WARPS == 1 || GRB_SOLVE_THREADS / (32 * GRB_GROUP_WARPS) <= 15) &&
              (GRB_GROUP_WARPS_LEAN == 1 || GRB_SOLVE_THREADS / (32 * GRB_GROUP_WARPS_LEAN) <= 15), "named barriers 1..15");

// barrier + OR-reduction of a predicate over the group
template <int GW>
__device__ __forceinline__ bool group_any(int bar_id, bool pred)
{
  int out;
  asm volatile(
      "{\n\t.reg .pred p, q;\n\tsetp.ne.s32 q, %3, 0;\n\tbar.red.or.pred p, %1, %2, q;\n\tselp.s32 %0, 1, 0, p;\n\t}"
      : "=r"(out)
      : "r"(bar_id), "r"(32 * GW), "r"((int)pred)
      : "memory");
  return out != 0;
}

// Warp-cooperative fetch of one table segment for the 32 cells a warp owns.
//   table   first (value, difference) pair of the segment in row 0
//   ROWB    bytes per table row;  NCOLS columns in the segment
//   my_off  THIS lane's cell: byte offset of its row; my_t: its interpolation weight
//   dst0    this warp's lane-0 slot row + the segment's first slot
// G lanes serve one cell per pass (32/G cells per pass, G passes): lane g of a group loads the
// pairs of columns g, g+G, g+2G, ... - each load instruction of the group covers G
// consecutive 16-byte words - forms a + d * t and stores column c to slot c of that cell's row.
// All loads of a pass are independent of each other.
// (explicit load batches over passes, a group barrier only every 2nd/4th subcycle, a
// prefetch.global.L1 of the stage-4 chemistry row issued before stage 2 (-4 %), and passing
// (weight, offset) through a shared-memory mailbox instead of the three shuffles were all
// measured slower than this loop as the compiler schedules it.)
template <int G, int NCOLS, int ROWB, int STRIDE = GRB_SLOT_STRIDE>
__device__ __forceinline__ void coop_fetch(const double *table, int my_off, double my_t, double *dst0)
{
  static_assert(G == 2 || G == 4 || G == 8 || G == 16, "lanes per cell");
  const unsigned full = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const int g = lane & (G - 1), sub = lane / G;
  constexpr int CPP = 32 / G;                 // cells per pass
  constexpr int PER = (NCOLS + G - 1) / G;    // columns per lane
  const char *src = reinterpret_cast<const char *>(table) + 16 * g;
  double *dst = dst0 + sub * STRIDE + g;
#pragma unroll
  for (int pass = 0; pass < G; pass++) {
    const int cell = pass * CPP + sub;
    const int off = __shfl_sync(full, my_off, cell);
    const double t = __shfl_sync(full, my_t, cell);
#pragma unroll
    for (int k = 0; k < PER; k++) {
      if ((k + 1) * G <= NCOLS || g + k * G < NCOLS) {
        const double2 p = __ldg(reinterpret_cast<const double2 *>(src + off + 16 * k * G));
        dst[pass * CPP * STRIDE + k * G] = p.x + p.y * t;
      }
    }
  }
}

// cool_multi_time_g.F: one evaluation of the cooling rate, t_cool = max(p/(gamma-1), tiny)/edot.
// Same structure as one subcycle of k_solve up to the cooling sums: per-cell temperature and
// log-T bin, warp-cooperative coalesced fetch of the cooling-table row into shared-memory
// slots (per-thread gathers kept the L1 at ~3/4 of its wavefront rate), log/exp table in
// shared memory.  VAR: optional physics modules folded away at compile time, as for k_solve.
#ifndef GRB_CT_THREADS
#define GRB_CT_THREADS 64    /* 2-warp blocks, 14 per SM: +4.4 % over 128 (17.2 vs 16.5 Gcell/s, pc=3 compact) */
#endif
#define GRB_CT_STRIDE 25
template <int PC, int VAR>
__global__ void __launch_bounds__(GRB_CT_THREADS) k_cooling_time(const __grid_constant__ GrParams P0)
{
  __shared__ double ct_slots[GRB_CT_THREADS * GRB_CT_STRIDE];
  __shared__ double ct_fm[GRB_FM_TABLE_DOUBLES];
  // the block's copy of the log/exp table: loads issued here, stored to shared memory only
  // after the cell's field loads are in flight too (one memory round trip instead of two).
  // (8 blocks per SM at 64 registers, and resident blocks striding over tiles with an L2
  // prefetch one tile ahead, both measured slower than this one-tile-per-block form.)
  constexpr int FM_PER = (GRB_FM_TABLE_DOUBLES + GRB_CT_THREADS - 1) / GRB_CT_THREADS;
  double fm_part[FM_PER];
#pragma unroll
  for (int k = 0; k < FM_PER; k++) {
    const int i = threadIdx.x + k * GRB_CT_THREADS;
    fm_part[k] = (i < GRB_FM_TABLE_DOUBLES) ? __ldg(P0.fm_tab + i) : 0.0;
  }
  GrParams P = P0;
  P.fm_tab = ct_fm;
  variant_flags<VAR>(P);
  const bool anydust = (P.idust > 0) || (P.idustall > 0) || (P.idustrec > 0);
  double *const my_slots = ct_slots + threadIdx.x * GRB_CT_STRIDE;
  double *const warp_slots = ct_slots + (threadIdx.x & ~31u) * GRB_CT_STRIDE;

  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const bool valid = t < P.nactive;
  Cell c;
  Thermo th;
  size_t x = 0;
  double tgasold = 0.0, logtem = 0.0, tdef = 0.0;
  int off = 0;
  if (valid) {
    int i, j, k;
    x = cell_index(P, t, i, j, k);
    load_cell<PC, true>(P, x, c, false);
  }
#pragma unroll
  for (int k = 0; k < FM_PER; k++) {
    const int i = threadIdx.x + k * GRB_CT_THREADS;
    if (i < GRB_FM_TABLE_DOUBLES) ct_fm[i] = fm_part[k];
  }
  __syncthreads();
  if (valid) {
    logtem = thermo_head<PC>(P, c, true, tgasold, th, false);
    if ((PC > 0) || (P.idustall > 0)) {           // cool1d_multi_g.F:349-357
      const TIndex ti = t_index(P, logtem);
      off = (ti.idx - 1) * (NCOOL * 16);
      tdef = ti.tdef;
    }
  }
  if (PC > 1 || anydust) coop_fetch<8, NCOOL, NCOOL * 16, GRB_CT_STRIDE>(P.cool, off, tdef, warp_slots);
  else if (PC == 1) coop_fetch<4, L_BREM + 1, NCOOL * 16, GRB_CT_STRIDE>(P.cool, off, tdef, warp_slots);
  __syncwarp();
  if (valid) {
    cool_sums<PC>(P, c, logtem, CoolSlots{my_slots}, tgasold, th);
    const double energy = dmax(th.p2d * P.inv_gm1, GRB_TINY);
    P.out[x] = ddiv_s(energy, th.edot);
  }
}

// ---- host-buffer pipeline hooks (GrParams::chunk_*; api.cu solve_pipelined) -----------
// They are a template switch of the kernel, not a run-time one: merely present in the
// subcycle loop they cost the device-resident path 2 % (measured), so that path runs
// instantiations without them.
// Per-lane state (the chunk of the cell in hand, the number
// of chunks last seen ready) lives in the spare 41st slot of the lane's shared-memory row,
// not in registers: the subcycle loop has none to give.
// Wait until `chunk` has arrived: *count is the number of chunks whose copies are complete,
// written by the copy stream behind each chunk's field copies.  Returns the count seen, or 0
// after 20 s without progress so that a host-side failure cannot wedge the GPU.
__device__ __forceinline__ unsigned chunk_wait(const unsigned *count, unsigned chunk)
{
  unsigned v;
  unsigned long long t0 = 0;
  for (;;) {
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(count) : "memory");
    if (v > chunk) return v;
    unsigned long long now;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
    if (t0 == 0) t0 = now;
    else if (now - t0 > 20000000000ull) return 0u;
    __nanosleep(500);
  }
}

// Finished cells are counted per lane (lane_state[0] = chunk << 16 | finished cells of that
// chunk not yet reported) and reported when the lane moves on to another chunk or leaves:
// the report needs a fence behind the cell's stores, and a fence per finished cell sat on the
// lockstep group's critical path (+4 % kernel time, measured).  The lane that completes the
// chunk's count tells the host.
__device__ __forceinline__ void chunk_report(const GrParams &P, unsigned state)
{
  const unsigned chunk = state >> 16, n = state & 0xffffu;
  if (n == 0u) return;
  const long long left = P.nactive - ((long long)chunk << P.chunk_shift);
  const long long size = 1ll << P.chunk_shift;
  const unsigned total = (unsigned)(left < size ? left : size);
  __threadfence();                       // the cells' stores before the count
  if (atomicAdd(P.chunk_done + chunk, n) + n == total) {
    __threadfence_system();
    *reinterpret_cast<volatile unsigned *>(P.chunk_done_host + chunk) = 1u;
  }
}

template <int PC, bool PIPE, int GW>
__device__ __forceinline__ void solve_cells(const GrParams &P)
{
  extern __shared__ double slot_rows[];     // [GRB_SOLVE_THREADS][GRB_SLOT_STRIDE]
  double *const my_slots = slot_rows + (size_t)threadIdx.x * GRB_SLOT_STRIDE;
  double *const warp_slots = slot_rows + (size_t)(threadIdx.x & ~31u) * GRB_SLOT_STRIDE;
  const Rates r{my_slots};
  const unsigned full = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const int bar_id = 1 + (int)(threadIdx.x >> 5) / GW;
  const bool anydust = (P.idust > 0) || (P.idustall > 0) || (P.idustrec > 0);
  Cell c;
  SubState s;
  size_t x = 0;
  bool active = false, exhausted = false;
  unsigned *const lane_state = reinterpret_cast<unsigned *>(my_slots + NSLOT);   // [0] chunk | unreported, [1] chunks seen ready
  if (PIPE) { lane_state[0] = 0u; lane_state[1] = 0u; }
  unsigned long long my_sub = 0, my_max = 0, my_itmax = 0, my_nan = 0;

  for (;;) {
    // ---- refill idle lanes from the global queue ----
    const unsigned need = __ballot_sync(full, !active && !exhausted);
    if (need) {
      unsigned long long base = 0;
      const int leader = __ffs(need) - 1;
      if (lane == leader) base = atomicAdd(P.work_counter, (unsigned long long)__popc(need));
      base = __shfl_sync(full, base, leader);
      if (!active && !exhausted) {
        const long long t = (long long)(base + __popc(need & ((1u << lane) - 1)));
        if (t < P.nactive) {
          int ci, cj, ck;
          x = cell_index(P, t, ci, cj, ck);
          bool ready = true;
          if (PIPE) {
            const unsigned chunk = (unsigned)(t >> P.chunk_shift), state = lane_state[0];
            if ((state >> 16) != chunk || (state & 0xffffu) == 0xffffu) {
              chunk_report(P, state);
              lane_state[0] = chunk << 16;
            }
            if (chunk >= lane_state[1]) ready = (lane_state[1] = chunk_wait(P.chunk_ready, chunk)) != 0u;
          }
          if (ready) {
            bool run = true;
            if (P.iradtrans == 1 && P.iradcoupled == 1) {   // coupled-RT masks (:418-439)
              const bool lit = P.f[F_KPHHI][x] > 0;
              run = (P.iradstep == 1) ? lit : !lit;
            }
            load_cell<PC, false>(P, x, c, true);
            s.ttot = 0.0; s.tgasold = 0.0; s.dedot_prev = 0.0; s.HIdot_prev = 0.0; s.iter = 1;
            if (run) active = true;
            else {                            // masked cell: floors + consistency only
              store_cell<PC>(P, x, c);
              if (PIPE) lane_state[0] += 1u;
            }
          } else {                            // the host never delivered this chunk
            atomicMax(P.diag + 5, 1ull);
            exhausted = true;
          }
        } else {
          exhausted = true;
          if (PIPE) {                         // nothing more to come: report now, not at group exit
            chunk_report(P, lane_state[0]);
            lane_state[0] &= 0xffff0000u;
          }
        }
      }
    }
    // after the refill every lane either owns a cell or has seen the end of the queue
    if (GW == 1 ? !__any_sync(full, active) : !group_any<GW>(bar_id, active)) break;

    // ---- one subcycle on every busy lane ----
    // stage 1 (per cell): temperature and the two log-T bins.  Idle lanes point at bin 1.
    Thermo th;
    TIndex tb;                       // bin of ln(tgas): chemistry rows
    tb.idx = 1; tb.tdef = 0.0; tb.t1 = tb.t2 = 0.0;
    int offA = 0, offB = 0;          // byte offsets of the lower rows (cooling, chemistry)
    double tdefA = 0.0, logtem = 0.0;
    if (active) {
      load_extras<PC>(P, x, c);
      logtem = thermo_head<PC>(P, c, s.iter == 1, s.tgasold, th, true);
      if ((PC > 0) || (P.idustall > 0)) {           // cool1d_multi_g.F:349-357
        const TIndex ta = t_index(P, logtem);
        offA = (ta.idx - 1) * (NCOOL * 16);
        tdefA = ta.tdef;
      }
      if (PC > 0) {                                 // solve_rate_cool_g.F:1211-1222
        double lt = dmax(th.logtgas, P.logtem0);
        lt = dmin(lt, P.logtem9);
        tb = t_index(P, lt);
        offB = (tb.idx - 1) * (NCHEM * 16);
      }
    }
    // stage 2 (warp): cooling rows and the k13dd coefficients -> slots
    if (PC > 1 || anydust) coop_fetch<GRB_G_COOL, NCOOL, NCOOL * 16>(P.cool, offA, tdefA, warp_slots + SLA_COOL);
    else if (PC == 1) coop_fetch<4, L_BREM + 1, NCOOL * 16>(P.cool, offA, tdefA, warp_slots + SLA_COOL);
    if (PC > 1 && P.ithreebody == 0)
      coop_fetch<GRB_G_DD, 14, NCHEM * 16>(P.chem + 2 * C_K13DD0, offB, tb.tdef, warp_slots + SLA_DD);
    __syncwarp();
    // stage 3 (per cell): cooling / heating sums, density-dependent k13
    double k13dd = GRB_TINY8;
    if (active) {
      cool_sums<PC>(P, c, logtem, CoolSlots{my_slots}, s.tgasold, th);
      if (PC > 1 && P.ithreebody == 0) k13dd = k13_martin(P, c, th, my_slots);
    }
    __syncwarp();
    if (PC > 0) {
      // stage 4 (warp): chemistry rows -> slots
      if (PC == 1) coop_fetch<4, chem_cols(1), NCHEM * 16>(P.chem, offB, tb.tdef, warp_slots);
      else coop_fetch<GRB_G_MAIN, chem_cols(PC), NCHEM * 16>(P.chem, offB, tb.tdef, warp_slots);
      __syncwarp();
      // stage 5 (per cell): shielding, H2 on dust
      if (active) subcycle_lookup<PC>(P, c, th, x, tb, k13dd, r);
    }
    if (active) {
      const bool done = subcycle_finish<PC>(P, c, s, th, tb, r);
      my_sub++;
      if (done || s.iter >= P.itmax) {
        if (!done) my_itmax++;
        if (s.iter > (int)my_max) my_max = s.iter;
        if (!(c.e == c.e) || !(c.HI == c.HI)) my_nan++;
        store_cell<PC>(P, x, c);
        if (PIPE) lane_state[0] += 1u;
        active = false;
      } else {
        s.iter++;
      }
    }
    __syncwarp();
  }

  if (PIPE) chunk_report(P, lane_state[0]);

  // ---- diagnostics: warp reduce, one atomic per warp ----
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    my_sub += __shfl_xor_sync(full, my_sub, o);
    my_itmax += __shfl_xor_sync(full, my_itmax, o);
    my_nan += __shfl_xor_sync(full, my_nan, o);
    const unsigned long long m = __shfl_xor_sync(full, my_max, o);
    my_max = m > my_max ? m : my_max;
  }
  if (lane == 0) {
    atomicAdd(P.diag + 0, my_sub);
    atomicMax(P.diag + 1, my_max);
    if (my_itmax) atomicAdd(P.diag + 2, my_itmax);
    if (my_nan) atomicAdd(P.diag + 3, my_nan);
  }
}

// Three instantiations per network.  The executed path of a subcycle is far larger than the
// SM's instruction cache, so what is NOT executed matters too: with every optional physics
// module compiled in, the hot instructions of the common configuration are scattered over
// 260 KB of SASS.  A variant folds the switches of modules it omits to "off" at compile time
// (a by-value copy of the parameters whose flags are overwritten with constants; the
// compiler propagates them and drops the dead code: 16.4k -> 5.6k instructions, +32 %
// cells/s at config C4):
//   VAR_GENERAL      everything, all switches read at run time
//   VAR_LEAN         network + metal cooling + UV background (configs C1, C2, C4)
//   VAR_DUST_SHIELD  VAR_LEAN + H2 formation on dust (h2_on_dust, no dust field) + UVB
//                    self-shielding, any method (config C3)
// grb_solve_variant_for() is the host-side choice; anything else runs VAR_GENERAL.
enum { VAR_GENERAL = 0, VAR_LEAN = 1, VAR_DUST_SHIELD = 2, VAR_RANK45 = 3 };

template <int VAR>
__host__ __device__ inline void variant_flags(GrParams &P)
{
  // VAR_RANK45 = VAR_GENERAL + old-style metal tables of rank 4/5: their 16- and 32-corner
  // interpolation (three lookups per subcycle) costs every general run up to 20 % through
  // register pressure when it is merely compiled in, so it gets its own instantiation
  if (VAR == VAR_RANK45) return;
  P.iold45 = 0;
  if (VAR == VAR_GENERAL) return;
  P.idustall = 0; P.idustfield = 0; P.idustrec = 0; P.igammah = 0; P.ih2optical = 0;
  P.iciecool = 0; P.ithreebody = 0; P.iH2shield = 0; P.iradtrans = 0; P.iradcoupled = 0;
  P.iVheat = 0; P.iMheat = 0; P.iisrffield = 0; P.iH2shieldcustom = 0; P.clnew = 1;
  P.met.fast = 1; P.pri.fast = 1;
  if (VAR == VAR_LEAN) { P.idust = 0; P.iradshield = 0; }
}

template <int PC, int VAR, bool PIPE>
__global__ void __launch_bounds__(GRB_SOLVE_THREADS, GRB_SOLVE_MINBLOCKS) k_solve(const __grid_constant__ GrParams P0)
{
  // the 2.5 KB log/exp table goes to shared memory behind the slot rows; the by-value copy of
  // the parameters is what lets the compiler fold the variant's switches and the table pointer
  extern __shared__ double slot_rows[];
  double *const fm = slot_rows + (size_t)GRB_SOLVE_THREADS * GRB_SLOT_STRIDE;
  for (int i = threadIdx.x; i < GRB_FM_TABLE_DOUBLES; i += GRB_SOLVE_THREADS) fm[i] = P0.fm_tab[i];
  __syncthreads();
  GrParams P = P0;
  P.fm_tab = fm;
  variant_flags<VAR>(P);
  solve_cells<PC, PIPE,
              (VAR != VAR_LEAN) ? ((PC <= 1) ? GRB_GROUP_WARPS : GRB_GROUP_WARPS_BIG)
                                : ((PC <= 1) ? GRB_GROUP_WARPS_LEAN_SMALL : GRB_GROUP_WARPS_LEAN)>(P);
}

// ------------------------------------------------------------------------------
// FP64 FMA-chain microbenchmark: the measured non-tensor FP64 peak that the solve
// kernel's roofline fraction is quoted against.
// 8 independent chains per thread, 2 flop per DFMA.
// ------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_fp64_peak(double *sink, int iters, double a, double b)
{
  double x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
  for (int i = 0; i < iters; i++) {
#pragma unroll
    for (int u = 0; u < 8; u++) {
      x0 = fma(x0, a, b); x1 = fma(x1, a, b); x2 = fma(x2, a, b); x3 = fma(x3, a, b);
      x4 = fma(x4, a, b); x5 = fma(x5, a, b); x6 = fma(x6, a, b); x7 = fma(x7, a, b);
    }
  }
  const double r = ((x0 + x1) + (x2 + x3)) + ((x4 + x5) + (x6 + x7));
  if (r == 123.456) sink[0] = r;      // never true; keeps the chains alive
}

// ------------------------------------------------------------------------------
// host-callable launchers (declared in gr_launch.h)
// ------------------------------------------------------------------------------
#define GRB_SOLVE_SMEM (((size_t)GRB_SLOT_STRIDE * GRB_SOLVE_THREADS + GRB_FM_TABLE_DOUBLES) * sizeof(double))
static inline unsigned blocks_for(long long n, int threads) { return (unsigned)((n + threads - 1) / threads); }

}  // namespace grb

using namespace grb;

#define DISPATCH_PC(pc, CALL)                 \
  switch (pc) {                               \
    case 0: { CALL(0); break; }               \
    case 1: { CALL(1); break; }               \
    case 2: { CALL(2); break; }               \
    default: { CALL(3); break; }              \
  }

void grb_launch_thermo(const GrParams &P, int mode, cudaStream_t st)
{
  const unsigned nb = blocks_for(P.nactive, 256);
#define CALL(PCV)                                                              \
  if (mode == 0) k_thermo<PCV, 0><<<nb, 256, 0, st>>>(P);                      \
  else if (mode == 1) k_thermo<PCV, 1><<<nb, 256, 0, st>>>(P);                 \
  else k_thermo<PCV, 2><<<nb, 256, 0, st>>>(P);
  DISPATCH_PC(P.ispecies, CALL)
#undef CALL
}

void grb_launch_fill(const GrParams &P, double value, cudaStream_t st)
{
  k_fill<<<blocks_for(P.nactive, 256), 256, 0, st>>>(P, value);
}

void grb_launch_temperature_table(const GrParams &P, cudaStream_t st)
{
  k_temperature_table<<<blocks_for(P.nactive, 256), 256, 0, st>>>(P);
}

void grb_launch_cooling_time(const GrParams &P, cudaStream_t st)
{
  const unsigned nb = blocks_for(P.nactive, GRB_CT_THREADS);
  const int var = grb_solve_variant_for(P);
#define CALL(PCV)                                                                      \
  if (var == VAR_LEAN) k_cooling_time<PCV, VAR_LEAN><<<nb, GRB_CT_THREADS, 0, st>>>(P);           \
  else if (var == VAR_DUST_SHIELD) k_cooling_time<PCV, VAR_DUST_SHIELD><<<nb, GRB_CT_THREADS, 0, st>>>(P); \
  else if (var == VAR_RANK45) k_cooling_time<PCV, VAR_RANK45><<<nb, GRB_CT_THREADS, 0, st>>>(P);            \
  else k_cooling_time<PCV, VAR_GENERAL><<<nb, GRB_CT_THREADS, 0, st>>>(P);
  DISPATCH_PC(P.ispecies, CALL)
#undef CALL
}

void grb_launch_dust_temperature(const GrParams &P, cudaStream_t st)
{
  const unsigned nb = blocks_for(P.nactive, 128);
#define CALL(PCV) k_dust_temperature<PCV><<<nb, 128, 0, st>>>(P);
  DISPATCH_PC(P.ispecies, CALL)
#undef CALL
}

int grb_solve_variant = 0;

// which instantiation serves this parameter set (VAR_* above)
int grb_solve_variant_for(const GrParams &P)
{
  if (P.iold45 != 0) return VAR_RANK45;
  if (grb_solve_variant != 0) return VAR_GENERAL;
  // "off" is <= 0 for every switch (unset options stay at -1, e.g. dust_recombination_cooling)
  const bool off = P.idustall <= 0 && P.idustfield <= 0 && P.idustrec <= 0 && P.igammah <= 0 &&
                   P.ih2optical <= 0 && P.iciecool <= 0 && P.iH2shield <= 0 &&
                   P.iradtrans <= 0 && P.iVheat <= 0 && P.iMheat <= 0 && P.iisrffield <= 0 && P.iH2shieldcustom <= 0;
  const bool ok = off && (P.ispecies < 2 || P.ithreebody == 0) && (P.imcool == 0 || (P.clnew == 1 && P.met.fast == 1)) &&
                  (P.ispecies > 0 || P.pri.fast == 1);
  if (!ok) return VAR_GENERAL;
  if (P.idust <= 0 && P.iradshield <= 0) return VAR_LEAN;
  return (P.ispecies > 0) ? VAR_DUST_SHIELD : VAR_GENERAL;
}

// persistent launch: `blocks` CTAs of GRB_SOLVE_THREADS threads (host picks a multiple of
// the SM count); P.work_counter must be zero
void grb_launch_solve(const GrParams &P, int blocks, cudaStream_t st)
{
  const int var = grb_solve_variant_for(P);
#define LAUNCH1(PCV, V, PIPE)                                                                                  \
  {                                                                                                              \
    cudaFuncSetAttribute(k_solve<PCV, V, PIPE>, cudaFuncAttributeMaxDynamicSharedMemorySize, GRB_SOLVE_SMEM);    \
    k_solve<PCV, V, PIPE><<<blocks, GRB_SOLVE_THREADS, GRB_SOLVE_SMEM, st>>>(P);                                 \
  }
#define LAUNCH(PCV, V)                                             \
  {                                                                \
    if (P.chunk_shift > 0) LAUNCH1(PCV, V, true)                   \
    else LAUNCH1(PCV, V, false)                                    \
  }
#define CALL(PCV)                                                   \
  if (var == VAR_LEAN) LAUNCH(PCV, VAR_LEAN)                        \
  else if (var == VAR_DUST_SHIELD) LAUNCH(PCV, VAR_DUST_SHIELD)     \
  else if (var == VAR_RANK45) LAUNCH(PCV, VAR_RANK45)               \
  else LAUNCH(PCV, VAR_GENERAL)
  DISPATCH_PC(P.ispecies, CALL)
#undef CALL
#undef LAUNCH
#undef LAUNCH1
}

int grb_solve_max_blocks_per_sm(int pc)
{
  int n = 0;
#define CALL(PCV)                                                                                           \
  cudaFuncSetAttribute(k_solve<PCV, VAR_GENERAL, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, GRB_SOLVE_SMEM);   \
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, k_solve<PCV, VAR_GENERAL, true>, GRB_SOLVE_THREADS, GRB_SOLVE_SMEM);
  DISPATCH_PC(pc, CALL)
#undef CALL
  return n;
}

// returns flops issued by one launch
double grb_launch_fp64_peak(double *sink, int blocks, int iters, cudaStream_t st)
{
  k_fp64_peak<<<blocks, 256, 0, st>>>(sink, iters, 0.999999, 1e-9);
  return (double)blocks * 256.0 * (double)iters * 64.0 * 2.0;
}
