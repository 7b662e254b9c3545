// grackle-b200: the sm_100a kernels.  One thread per active cell, whole cell state in
// registers, SoA fields read/written exactly once per call, comoving<->proper scaling
// and species floors/renormalisation fused into the same pass.
//
//   k_pressure / k_temperature / k_gamma   calculate_{pressure,temperature,gamma}.c   HBM-bound
//   k_temperature_table                    calc_temp_cloudy_g.F (pc = 0)              HBM + L2 gathers
//   k_cooling_time                         cool_multi_time_g.F                        FP64 + HBM
//   k_dust_temperature                     calc_tdust_3d_g.F                          FP64 (Newton)
//   k_solve (solve_kernel.cuh, solve_pc*.cu) solve_rate_cool_g.F (whole file)        FP64 pipe / LSU data pipe
//
// No tensor cores: there is no dense contraction anywhere on this path (north_star).
#include <cuda_runtime.h>

#include "cell_io.cuh"

namespace grb {

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
        // lean exp and divide (fastmath.cuh; 0 < xx < 10, so ex - 1 > 0): the library exp and two IEEE
        // divisions here were a quarter of the instructions of every warp that holds one H2-rich cell
        const double ex = dexp(xx);
        GammaH2Inverse = 0.5 * (5 + ddiv_s(2.0 * xx * xx * ex, (ex - 1.0) * (ex - 1.0)));
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
      const double ex = dexp(xx);
      GammaH2Inverse = 0.5 * (5 + ddiv_s(2.0 * xx * xx * ex, (ex - 1.0) * (ex - 1.0)));
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

// cool_multi_time_g.F: one evaluation of the cooling rate, t_cool = max(p/(gamma-1), tiny)/edot.
// Same structure as one subcycle of k_solve up to the cooling sums: per-cell temperature and
// log-T bin, warp-cooperative coalesced fetch of the cooling-table row into shared-memory
// slots (per-thread gathers kept the L1 at ~3/4 of its wavefront rate), log/exp table in
// shared memory.  VAR: optional physics modules folded away at compile time, as for k_solve.
// Block shape (pc = 3 compact variant, 67 M cells, Gcell/s): round 1 shipped 64-thread blocks, 14 per
// SM by shared memory (17.2).  Round 2: the compact variant keeps only the 20 cooling columns it
// reads (13.8 -> 11.7 KB of slots per 64 threads) and its occupancy is pinned through
// __launch_bounds__: 64 threads x 12 / 14 / 16 blocks (80 / 71 / 64 registers) 16.6 / 17.9 / 16.1,
// 128 threads x 7 / 8 blocks (71 / 64 registers) 17.8 / 19.05.  The other variants need more
// registers and keep the compiler's choice.
#ifndef GRB_CT_THREADS
#define GRB_CT_THREADS 128
#endif
#ifndef GRB_CT_MINBLOCKS
#define GRB_CT_MINBLOCKS (1024 / GRB_CT_THREADS)   /* compact variant: 32 warps per SM at 64 registers */
#endif
__host__ __device__ constexpr int ct_minblocks(int var) { return var == VAR_LEAN ? GRB_CT_MINBLOCKS : (640 / GRB_CT_THREADS > 0 ? 640 / GRB_CT_THREADS : 1); }
template <int PC, int VAR>
__global__ void __launch_bounds__(GRB_CT_THREADS, ct_minblocks(VAR)) k_cooling_time(const __grid_constant__ GrParams P0)
{
  constexpr int G = coop_lanes(PC);
  constexpr int CS = slot_col_stride(GRB_CT_THREADS, G);      // column-major slots (network.cuh)
  constexpr int NCA = cool_cols(PC, VAR == VAR_LEAN);         // the compact variant reads no dust / CIE columns
  __shared__ double ct_slots[NCA * CS];
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
  double *const my_slots = ct_slots + threadIdx.x;
  double *const warp_slots = ct_slots + (threadIdx.x & ~31u);

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
    load_cell<PC, true, false>(P, x, c, false);
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
  if (PC > 1 || anydust) coop_fetch<G, NCA, NCOOL * 16, CS>(P.cool, off, tdef, warp_slots);
  else if (PC == 1) coop_fetch<G, L_BREM + 1, NCOOL * 16, CS>(P.cool, off, tdef, warp_slots);
  __syncwarp();
  if (valid) {
    cool_sums<PC>(P, c, logtem, CoolSlots<CS>{my_slots}, tgasold, th);
    const double energy = dmax(th.p2d * P.inv_gm1, GRB_TINY);
    P.out[x] = ddiv_s(energy, th.edot);
  }
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
  else if (var == VAR_DUST_ALL) k_cooling_time<PCV, VAR_DUST_ALL><<<nb, GRB_CT_THREADS, 0, st>>>(P);       \
  else if (var == VAR_COMMON) k_cooling_time<PCV, VAR_COMMON><<<nb, GRB_CT_THREADS, 0, st>>>(P);           \
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

// which instantiation serves this parameter set (VAR_* in cell_io.cuh)
int grb_solve_variant_for(const GrParams &P)
{
  if (P.iold45 != 0) return VAR_RANK45;
  if (grb_solve_variant != 0) return VAR_GENERAL;
  // "off" is <= 0 for every switch (unset options stay at -1, e.g. dust_recombination_cooling)
  // no field-array inputs, new-style tables
  const bool base = P.idustfield <= 0 && P.iradtrans <= 0 && P.iVheat <= 0 && P.iMheat <= 0 && P.iisrffield <= 0 &&
                    P.iH2shieldcustom <= 0 && (P.imcool == 0 || (P.clnew == 1 && P.met.fast == 1)) &&
                    (P.ispecies > 0 || P.pri.fast == 1);
  if (!base) return VAR_GENERAL;
  const bool nodustall = P.idustall <= 0 && P.idustrec <= 0 && P.igammah <= 0;
  const bool off = P.ih2optical <= 0 && P.iciecool <= 0 && P.iH2shield <= 0 && (P.ispecies < 2 || P.ithreebody == 0);
  // the H2 networks with the collapse / galaxy-run switches on: Jeans-length H2 shielding, three-body rates,
  // CIE cooling, optical-depth fudge (VAR_COMMON)
  if (!off) return (nodustall && P.ispecies > 1 && (P.iH2shield <= 0 || P.iH2shield == 3)) ? VAR_COMMON : VAR_GENERAL;
  if (nodustall && P.idust <= 0 && P.iradshield <= 0) return VAR_LEAN;
  if (nodustall) return (P.ispecies > 0) ? VAR_DUST_SHIELD : VAR_GENERAL;
  // dust_chemistry = 1 with the settings it implies (initialize_chemistry_data.c:71-100), no UVB shielding
  if (P.idustall == 1 && P.idustrec == 1 && P.igammah == 2 && (P.ispecies < 2 || P.idust == 1) && P.iradshield <= 0 &&
      P.ispecies > 0 && P.imetal == 1)
    return VAR_DUST_ALL;
  return VAR_GENERAL;
}

// persistent launch: `blocks` CTAs of GRB_SOLVE_THREADS threads (host picks a multiple of
// the SM count); P.work_counter must be zero.  The instantiations live in solve_pc*.cu.
void grb_launch_solve_pc0(const GrParams &P, int var, int blocks, cudaStream_t st);
void grb_launch_solve_pc1(const GrParams &P, int var, int blocks, cudaStream_t st);
void grb_launch_solve_pc2(const GrParams &P, int var, int blocks, cudaStream_t st);
void grb_launch_solve_pc3(const GrParams &P, int var, int blocks, cudaStream_t st);
int grb_solve_max_blocks_pc0();
int grb_solve_max_blocks_pc1();
int grb_solve_max_blocks_pc2();
int grb_solve_max_blocks_pc3();

void grb_launch_solve(const GrParams &P, int blocks, cudaStream_t st)
{
  const int var = grb_solve_variant_for(P);
  switch (P.ispecies) {
    case 0: grb_launch_solve_pc0(P, var, blocks, st); break;
    case 1: grb_launch_solve_pc1(P, var, blocks, st); break;
    case 2: grb_launch_solve_pc2(P, var, blocks, st); break;
    default: grb_launch_solve_pc3(P, var, blocks, st); break;
  }
}

int grb_solve_max_blocks_per_sm(int pc)
{
  switch (pc) {
    case 0: return grb_solve_max_blocks_pc0();
    case 1: return grb_solve_max_blocks_pc1();
    case 2: return grb_solve_max_blocks_pc2();
    default: return grb_solve_max_blocks_pc3();
  }
}

// returns flops issued by one launch
double grb_launch_fp64_peak(double *sink, int blocks, int iters, cudaStream_t st)
{
  k_fp64_peak<<<blocks, 256, 0, st>>>(sink, iters, 0.999999, 1e-9);
  return (double)blocks * 256.0 * (double)iters * 64.0 * 2.0;
}
