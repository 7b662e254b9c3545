// grackle-b200: what the kernels share - active-cell indexing, one cell's load / store with the
// comoving <-> proper scaling and species floors fused in, the warp-cooperative table fetch and
// the compile-time physics variants.
#ifndef GRB200_CELL_IO_CUH
#define GRB200_CELL_IO_CUH

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

// Per-cell source terms that only the solve reads (radiative-transfer ionisation rates, the
// H2 shielding length / custom factor).  The cooling-time kernel never calls this: those
// field pointers are not bound for it (api.cu: needed_fields).
template <int PC>
__device__ __forceinline__ void load_solve_sources(const GrParams &P, size_t x, Cell &c)
{
  if (P.iradtrans == 1) {
    c.kphHI = P.f[F_KPHHI][x]; c.kphHeI = P.f[F_KPHHEI][x]; c.kphHeII = P.f[F_KPHHEII][x];
    c.kdissH2I = (PC > 1) ? P.f[F_KDISSH2I][x] : 0.0;
  } else {
    c.kphHI = c.kphHeI = c.kphHeII = c.kdissH2I = 0.0;
  }
  c.xH2shield = (PC > 1 && P.iH2shield == 2) ? P.f[F_XH2SHIELD][x] : 0.0;
  c.fshield = (PC > 1 && P.iH2shieldcustom > 0) ? P.f[F_FSHIELD_CUSTOM][x] : 1.0;
}

// ... and those the cooling rate itself reads (heating fields, ISRF, RT heating)
__device__ __forceinline__ void load_cool_sources(const GrParams &P, size_t x, Cell &c)
{
  c.vheat = (P.iVheat == 1) ? P.f[F_VHEAT][x] : 0.0;
  c.mheat = (P.iMheat == 1) ? P.f[F_MHEAT][x] : 0.0;
  c.isrf = (P.iisrffield > 0) ? P.f[F_ISRF][x] : 0.0;
  c.photogamma = (P.iradtrans == 1) ? P.f[F_PHOTOGAMMA][x] : 0.0;
}

// Load one cell, converting comoving -> proper in registers (scale_fields_g) and applying
// the species floors (ceiling_species_g) when `floors` is set (solve only).
template <int PC, bool SCALE_DIV, bool SOLVE>
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
  load_cool_sources(P, x, c);
  if (SOLVE) {
    load_solve_sources<PC>(P, x, c);
  } else {
    c.kphHI = c.kphHeI = c.kphHeII = c.kdissH2I = c.xH2shield = 0.0;
    c.fshield = 1.0;
  }
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

// Warp-cooperative fetch of one table segment for the 32 cells a warp owns.
//   table   first (value, difference) pair of the segment in row 0
//   ROWB    bytes per table row;  NCOLS columns in the segment;  CS slot column stride
//   my_off  THIS lane's cell: byte offset of its row; my_t: its interpolation weight
//   dst0    slot (first column of the segment, this warp's lane-0 thread)
// G lanes serve one cell per pass (32/G cells per pass, G passes): lane g of a group loads the
// pairs of columns g, g+G, g+2G, ... - each load instruction of the group covers G
// consecutive 16-byte words - forms a + d * t and stores column c of that cell.  With the
// column-major slots (network.cuh) the 16 lanes of a half-warp store to 16 different banks.
// All loads of a pass are independent of each other.
// (explicit load batches over passes, a group barrier only every 2nd/4th subcycle, a
// prefetch.global.L1 of the stage-4 chemistry row issued before stage 2 (-4 %), and passing
// (weight, offset) through a shared-memory mailbox instead of the three shuffles were all
// measured slower than this loop as the compiler schedules it.)
// Row offsets are non-negative and < 2^32: added as UNSIGNED 32-bit values so that the load can
// take the "64-bit base + 32-bit offset" address form instead of a sign extension and a 64-bit
// add per address (GRB_OFF_SIGNED=1 restores the signed form for comparison).
#ifndef GRB_OFF_SIGNED
#define GRB_OFF_SIGNED 0
#endif
#if GRB_OFF_SIGNED
#define GRB_OFF(x) (x)
#else
#define GRB_OFF(x) ((unsigned)(x))
#endif
template <int G, int NCOLS, int ROWB, int CS>
__device__ __forceinline__ void coop_fetch(const double *table, int my_off, double my_t, double *dst0)
{
  static_assert(G == 2 || G == 4 || G == 8 || G == 16, "lanes per cell");
  const unsigned full = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const int g = lane & (G - 1), sub = lane / G;
  constexpr int CPP = 32 / G;                 // cells per pass
  constexpr int PER = (NCOLS + G - 1) / G;    // columns per lane
  const char *src = reinterpret_cast<const char *>(table) + 16 * g;
  double *dst = dst0 + g * CS + sub;
#pragma unroll
  for (int pass = 0; pass < G; pass++) {
    const int cell = pass * CPP + sub;
    const int off = __shfl_sync(full, my_off, cell);
    const double t = __shfl_sync(full, my_t, cell);
#pragma unroll
    for (int k = 0; k < PER; k++) {
      if ((k + 1) * G <= NCOLS || g + k * G < NCOLS) {
        const double2 p = __ldg(reinterpret_cast<const double2 *>(src + GRB_OFF(off) + 16 * k * G));
        dst[pass * CPP + k * G * CS] = p.x + p.y * t;
      }
    }
  }
}

// Two segments in ONE pass loop (phase A of the H2 networks: NC1 cooling columns at row offset
// off1 / weight t1, then NC2 = 14 k13dd columns of the chemistry table at off2 / t2, stored to
// consecutive slot columns).  Written as one loop so that the loads of both segments of a pass -
// and of the next pass - are in flight together: as two separate coop_fetch loops the compiler
// serialised the second one (ncu, round 2: 13 fully exposed L2 round trips per subcycle, 17 % of
// the kernel's time).
// GRB_SKIP_SEG2 = 1: cells that will not read segment 2 this subcycle pass a negative offset and
// their lanes issue no load (fewer 128-byte lines per instruction).  MEASURED slower (B200, C4:
// 38.45 against 39.41 Mcell/s): the predicate costs more issue slots than the lines save.
#ifndef GRB_SKIP_SEG2
#define GRB_SKIP_SEG2 0
#endif
#ifndef GRB_FETCH_BATCH
#define GRB_FETCH_BATCH 2      /* passes whose loads are issued before the first of them is consumed */
#endif
template <int G, int NC1, int ROWB1, int NC2, int ROWB2, int CS>
__device__ __forceinline__ void coop_fetch2(const double *table1, int my_off1, double my_t1, const double *table2,
                                            int my_off2, double my_t2, double *dst0)
{
  const unsigned full = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const int g = lane & (G - 1), sub = lane / G;
  constexpr int CPP = 32 / G;
  constexpr int PER1 = (NC1 + G - 1) / G, PER2 = (NC2 + G - 1) / G;
  constexpr int NB = GRB_FETCH_BATCH;
  static_assert(G % NB == 0, "batch size");
  const char *src1 = reinterpret_cast<const char *>(table1) + 16 * g;
  const char *src2 = reinterpret_cast<const char *>(table2) + 16 * g;
  double *dst = dst0 + g * CS + sub;
#pragma unroll
  for (int pass0 = 0; pass0 < G; pass0 += NB) {
    double t1[NB], t2[NB];
    double2 p1[NB][PER1], p2[NB][PER2];
#pragma unroll
    for (int b = 0; b < NB; b++) {
      const int cell = (pass0 + b) * CPP + sub;
      const int off1 = __shfl_sync(full, my_off1, cell);
      const int off2 = __shfl_sync(full, my_off2, cell);
      t1[b] = __shfl_sync(full, my_t1, cell);
      t2[b] = __shfl_sync(full, my_t2, cell);
#pragma unroll
      for (int k = 0; k < PER1; k++)
        if ((k + 1) * G <= NC1 || g + k * G < NC1)
          p1[b][k] = __ldg(reinterpret_cast<const double2 *>(src1 + GRB_OFF(off1) + 16 * k * G));
      // off2 < 0: this cell does not read segment 2 this subcycle (k13_martin is only evaluated for
      // 500 K <= T < 1e6 K; idle lanes): its lanes issue no load, so the instruction touches fewer
      // 128-byte lines - the kernel is bound by the wavefronts of the load/store data pipe
#pragma unroll
      for (int k = 0; k < PER2; k++)
        if ((k + 1) * G <= NC2 || g + k * G < NC2) {
          p2[b][k] = make_double2(0.0, 0.0);
          if (!GRB_SKIP_SEG2 || off2 >= 0)
            p2[b][k] = __ldg(reinterpret_cast<const double2 *>(src2 + GRB_OFF(off2) + 16 * k * G));
        }
    }
#pragma unroll
    for (int b = 0; b < NB; b++) {
#pragma unroll
      for (int k = 0; k < PER1; k++)
        if ((k + 1) * G <= NC1 || g + k * G < NC1) dst[(pass0 + b) * CPP + k * G * CS] = p1[b][k].x + p1[b][k].y * t1[b];
#pragma unroll
      for (int k = 0; k < PER2; k++)
        if ((k + 1) * G <= NC2 || g + k * G < NC2)
          dst[(pass0 + b) * CPP + (NC1 + k * G) * CS] = p2[b][k].x + p2[b][k].y * t2[b];
    }
  }
}

// Compile-time physics variants.  The executed path of a subcycle is far larger than the
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
//   VAR_RANK45       VAR_GENERAL + old-style metal tables of rank 4/5
//   VAR_DUST_ALL     VAR_LEAN + the full dust model (dust_chemistry = 1: H2 on dust, gas-grain
//                    transfer, photo-electric heating, grain recombination cooling with the
//                    uniform interstellar radiation field; no dust / ISRF field arrays) - the
//                    parameter set of the reference's own examples (c_example.c:52-70,
//                    cxx_omp_example.C:160-180)
//   VAR_COMMON       VAR_DUST_SHIELD + the switches collapse / galaxy runs turn on, read at run time: the H2
//                    self-shielding length from the local Jeans length (H2_self_shielding = 3), any
//                    three_body_rate, CIE cooling and the H2 optical-depth fudge.  Still folded off: field-array
//                    inputs (radiative transfer, heating rates, dust / ISRF / shielding-length fields), the
//                    neighbour stencil, the full dust model, old-style tables.  9.1 k instructions against
//                    the general kernel's 13.7 k.
// grb_solve_variant_for() is the host-side choice; anything else runs VAR_GENERAL.
enum { VAR_GENERAL = 0, VAR_LEAN = 1, VAR_DUST_SHIELD = 2, VAR_RANK45 = 3, VAR_DUST_ALL = 4, VAR_COMMON = 5 };

template <int VAR>
__host__ __device__ inline void variant_flags(GrParams &P)
{
  // VAR_RANK45 = VAR_GENERAL + old-style metal tables of rank 4/5: their 16- and 32-corner
  // interpolation (three lookups per subcycle) costs every general run up to 20 % through
  // register pressure when it is merely compiled in, so it gets its own instantiation
  if (VAR == VAR_RANK45) return;
  P.iold45 = 0;
  if (VAR == VAR_GENERAL) return;
  P.idustfield = 0; P.iradtrans = 0; P.iradcoupled = 0;
  if (VAR == VAR_COMMON) {
    P.iH2shield = (P.iH2shield != 0) ? 3 : 0;      // the host only sends 0 or 3 here: lets the compiler drop options 1 and 2
  } else {
    P.ih2optical = 0; P.iciecool = 0; P.ithreebody = 0; P.iH2shield = 0;
  }
  P.iVheat = 0; P.iMheat = 0; P.iisrffield = 0; P.iH2shieldcustom = 0; P.clnew = 1;
  P.met.fast = 1; P.pri.fast = 1;
  if (VAR == VAR_DUST_ALL) {   // what dust_chemistry = 1 switches on (initialize_chemistry_data.c:71-100)
    P.idustall = 1; P.idust = 1; P.idustrec = 1; P.igammah = 2; P.iradshield = 0;
    return;
  }
  P.idustall = 0; P.idustrec = 0; P.igammah = 0;
  if (VAR == VAR_LEAN) { P.idust = 0; P.iradshield = 0; }
  if (VAR == VAR_COMMON) P.idust = (P.idust > 0) ? 1 : 0;
}
// variants whose physics switches are compile-time constants (network.cuh: FOLD)
__host__ __device__ constexpr bool variant_folds(int var) { return var != VAR_GENERAL && var != VAR_RANK45; }

}  // namespace grb
#endif
