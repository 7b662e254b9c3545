// grackle-b200: rate lookup, rate-limited time step and the backward-Euler species
// update for one cell (continuation of chem.cuh).
#ifndef GRB200_NETWORK_CUH
#define GRB200_NETWORK_CUH

#include "chem.cuh"
#include "gr_launch.h"

namespace grb {

// The interpolated rates of one cell-subcycle live in SHARED memory, COLUMN-major:
//     slot(col, thread) = slots[col * CS + thread],   CS = threads per CTA + 16 / G
// (G = lanes per cell of the cooperative fetch, kernels: coop_fetch).  A thread reads its own
// values at compile-time offsets col * CS from its base pointer; a warp's 32 reads of one slot
// are 32 consecutive doubles (2 wavefronts, the minimum for 256 bytes).  The slots are FILLED
// warp-cooperatively: G lanes read one cell's table row as consecutive 16-byte (value,
// difference) pairs - coalesced - interpolate, and store column c of that cell; the 16 lanes
// of a half-warp then write (G columns) x (16 / G cells): with the column stride CS = 16 / G
// (mod 16) these are 16 different 8-byte banks, so the stores are conflict-free as well.
// (Round 1 kept one row of 41 doubles per thread: conflict-free for the reads, but two lanes
// of every half-warp store collided - ncu: 296 shared-store wavefronts per warp-subcycle for
// 155 ideal, with the LSU data pipe at 76 % of its peak the kernel's tightest resource.)
//
// Slot map.  Phase A (before the cooling sums): the 24 cooling columns -> slots [0,24), the 14
// k13dd columns -> [24,38).  Phase B (before the species update) overwrites them: chemistry
// columns [0,32) -> slots [0,32), followed by the slots the lookup itself produces (only the
// variants with shielding / radiative transfer / dust write those; the compact variants read
// the call's constants instead).
#define GRB_NSLOTS 41
enum { SL_h2dust = GRB_NCHEM_MAIN, SL_k24s, SL_k25s, SL_k26s, SL_k28s, SL_k29s, SL_k30s, SL_k31s, NSLOT };
static_assert(NSLOT < GRB_NSLOTS, "slot columns (one spare: per-lane state of the host-buffer pipeline)");
enum { SLA_COOL = 0 };
// The compact variants (no dust, no CIE) use fewer slot columns: the cooling columns a network
// reads are the first 12 / 18 / 20 of the packed row, the 14 k13dd columns follow them directly,
// and phase B needs only the chemistry columns.  34 columns instead of 41 for pc = 3 (140 KB of
// shared memory instead of 169 KB): the SM's unified L1 / shared-memory array then leaves 92 KB
// of L1 for the table rows instead of 60 KB.
__host__ __device__ constexpr int cool_cols(int pc, bool compact)
{
  return !compact ? NCOOL : (pc == 1 ? L_BREM + 1 : (pc == 2 ? L_H2LTE + 1 : (pc == 3 ? L_HDLOW + 1 : NCOOL)));
}
__host__ __device__ constexpr int slot_cols_needed(int pc, bool compact, bool pipe)
{
  if (!compact) return GRB_NSLOTS;
  const int a = cool_cols(pc, true) + (pc > 1 ? 14 : 0);
  const int b = pc == 0 ? 0 : (pc == 1 ? C_k58 + 1 : (pc == 2 ? C_ncrd2 + 1 : C_k56 + 1));
  return (a > b ? a : b) + (pipe ? 1 : 0);     // + the per-lane state column of the host-buffer pipeline
}
static_assert(NCOOL + 14 <= GRB_NSLOTS, "phase A slots");

// chemistry columns the network PC needs every subcycle
__host__ __device__ constexpr int chem_cols(int pc) { return pc == 1 ? C_k58 + 1 : (pc == 2 ? C_ncrd2 + 1 : C_k56 + 1); }
// lanes per cell in the cooperative fetches of network PC, and the column stride that goes with it
__host__ __device__ constexpr int coop_lanes(int pc) { return pc == 1 ? 4 : 8; }
__host__ __device__ constexpr int slot_col_stride(int threads, int g) { return threads + 16 / g; }

// one thread's view of the slots
template <int CS>
struct Slots {
  double *base;    // &slots[0 * CS + thread]
  __device__ __forceinline__ double &operator[](int col) const { return base[col * CS]; }
};
#define RK(name) (r[C_##name])
#define RS(name) (r[SL_##name])
// cooling-table values read back from the slots (phase A)
template <int CS>
struct CoolSlots {
  const double *base;
  __device__ __forceinline__ double operator()(int col) const { return base[(SLA_COOL + col) * CS]; }
};

// neighbour densities for the Sobolev-like H2 shielding length (H2_self_shielding == 1)
struct Neigh { double d[6]; };

// ---------------------------------------------------------------------------------
// lookup_cool_rates1d_g (solve_rate_cool_g.F:1079-1737).  The plain log-T interpolations
// (:1225-1320) are done by coop_fetch; this is everything else, per cell.
// k13_martin: the density-dependent H2 collisional dissociation rate (Martin+96, :1709-1733)
// from the 14 interpolated coefficients `dd` (phase A slots).
// ---------------------------------------------------------------------------------
template <int DD0, class M, class SL>
__device__ __forceinline__ double k13_martin_m(M &m, double nh, const SL &slots)
{
#define DD(n) (slots[DD0 + (n)])
  // the four powers are independent: with the policy's unchecked cores they share one basic block
  const double p1 = m.pow(ddiv(nh, DD(4)), DD(6)), p2 = m.pow(ddiv(nh, DD(5)), DD(6));
  const double p3 = m.pow(ddiv(nh, DD(11)), DD(13)), p4 = m.pow(ddiv(nh, DD(12)), DD(13));
  double cid = DD(0) - ddiv(DD(1), 1.0 + p1) + DD(2) - ddiv(DD(3), 1.0 + p2);
  double dtn = DD(7) - ddiv(DD(8), 1.0 + p3) + DD(9) - ddiv(DD(10), 1.0 + p4);
  cid = dmax(m.exp10(cid), GRB_TINY8);
  dtn = dmax(m.exp10(dtn), GRB_TINY8);
  return dtn + cid;
#undef DD
}

template <int DD0, class SL>
__device__ inline double k13_martin(const GrParams &P, const Cell &c, const Thermo &th, const SL &slots)
{
  const double nh = dmin(c.HI * P.dom, 1.0e9);
  double k13 = GRB_TINY8;
  if (th.tgas >= 500.0 && th.tgas < 1.0e6) {
    FmFast f{P.fm_tab, false};
    k13 = k13_martin_m<DD0>(f, nh, slots);
    if (f.bad) {
      FmChecked e{P.fm_tab};
      k13 = k13_martin_m<DD0>(e, nh, slots);
    }
  }
  return k13;
}

// Which of the per-cell rate slots a parameter set actually produces.  FOLD = the kernel
// variant fixes its physics switches at compile time (every variant but the general ones):
// where a switch is off, the consumers then read the call's constant straight from the
// parameter block instead of a slot - no shared-memory store / load round trip for values
// that never change (8 stores and ~20 loads per cell-subcycle in the compact pc=3 kernel,
// whose tightest resource is the LSU data pipe).  The general variants (FOLD = false) keep
// every slot valid, as before.
__device__ __forceinline__ bool uvb_shielded(const GrParams &P) { return P.iradshield > 0; }
template <int PC>
__device__ __forceinline__ bool k31_per_cell(const GrParams &P)
{
  return PC > 1 && (P.iH2shield > 0 || P.iH2shieldcustom > 0 || P.iradtrans != 0);
}
#define Q_SLOT_UVB (!FOLD || uvb_shielded(P))
#define Q_k24s (Q_SLOT_UVB ? RS(k24s) : P.k24)
#define Q_k25s (Q_SLOT_UVB ? RS(k25s) : P.k25)
#define Q_k26s (Q_SLOT_UVB ? RS(k26s) : P.k26)
#define Q_k28s (Q_SLOT_UVB ? RS(k28s) : P.k28)
#define Q_k29s (Q_SLOT_UVB ? RS(k29s) : P.k29)
#define Q_k30s (Q_SLOT_UVB ? RS(k30s) : P.k30)
#define Q_k31s ((!FOLD || k31_per_cell<PC>(P)) ? RS(k31s) : P.k31)      /* read for PC > 1 only */
#define Q_h2dust (RS(h2dust))                                            /* read under anydust only */

// ti = the log-T bin of ln(tgas) (the one the chemistry rows were fetched at)
template <int PC, bool FOLD, class R>
__device__ inline void lookup_rates(const GrParams &P, const Cell &c, const Thermo &th,
                                    const Neigh &nb, const TIndex &ti, const R &r)
{
  const bool anydust = (P.idust > 0) || (P.idustall > 0);
  const double dom = P.dom;
  const double tdef = ti.tdef;
  if (!FOLD) {
    RS(k24s) = P.k24; RS(k25s) = P.k25; RS(k26s) = P.k26; RS(k28s) = P.k28; RS(k29s) = P.k29; RS(k30s) = P.k30;
    RS(k31s) = 0.0; RS(h2dust) = 0.0;
  }

  // H2 formation on grains: bilinear in (ln T, ln T_dust) (:1327-1378)
  if (anydust) {
    if (th.tdust > P.dtemend) {
      RS(h2dust) = GRB_TINY8;
    } else {
      double dl = DLOG(th.tdust);
      dl = dmax(dl, P.d_logtem0);
      dl = dmin(dl, P.d_logtem9);
      const long long di = imin(P.ndratec - 1, imax(1, (long long)((dl - P.d_logtem0) / P.d_dlogtem) + 1));
      const double d_t1 = (P.d_logtem0 + (di - 1) * P.d_dlogtem);
      const double d_t2 = (P.d_logtem0 + (di) * P.d_dlogtem);
      const double d_tdef = (dl - d_t1) / (d_t2 - d_t1);
      const double *c1 = P.h2dust + (size_t)(di - 1) * P.nratec + (ti.idx - 1);
      const double *c2 = c1 + P.nratec;
      const double dusti1 = ld(c1) + (ld(c1 + 1) - ld(c1)) * tdef;
      const double dusti2 = ld(c2) + (ld(c2 + 1) - ld(c2)) * tdef;
      RS(h2dust) = (dusti1 + (dusti2 - dusti1) * d_tdef) * th.dust2gas;
    }
  }

  if (PC > 1 && (!FOLD || k31_per_cell<PC>(P))) {     // H2 self-shielding, Wolcott-Green & Haiman (2019) (:1397-1490)
    double k31s = (P.iradtrans == 0) ? P.k31 : P.k31 + c.kdissH2I;
    if (P.iH2shield > 0) {
      double l_H2shield;
      if (P.iH2shield == 1) {
        double divrho = GRB_TINY;
#pragma unroll
        for (int n = 0; n < 6; n++) {
          const double dr = nb.d[n] - c.d;
          if (dr < 0.0) divrho = divrho + dr;
        }
        l_H2shield = dmin(P.dx_cgs * c.d / fabs(divrho), P.xbase1);
      } else if (P.iH2shield == 2) {
        l_H2shield = c.xH2shield * P.xbase1;
      } else if (P.iH2shield == 3) {
        l_H2shield = P.c_ljeans * sqrt(th.tgas / (c.d * th.mmw));
      } else {
        l_H2shield = 0.0;
      }
      const double N_H2 = dom * c.H2I * l_H2shield;
      double tuse = dmax(th.tgas, 1e2);
      tuse = dmin(tuse, 8e3);
      double nuse = c.d * dom / th.mmw;
      nuse = dmin(nuse, 1e7);
      const double aWG = (0.8711 * DLOG10(tuse) - 1.928) * DEXP(-0.2856 * DLOG10(nuse)) +
                         (-0.9639 * DLOG10(tuse) + 3.892);
      const double x = 2.0e-15 * N_H2;
      const double b_doppler = 1e-5 * sqrt(2.0 * GRB_KBOLTZ * th.tgas / GRB_MASS_H);
      double f_shield = 0.965 / POW_POS(1.0 + x / b_doppler, aWG) +
                        0.035 * DEXP(-8.5e-4 * sqrt(1.0 + x)) / sqrt(1.0 + x);
      f_shield = dmin(f_shield, 1.0);
      k31s = f_shield * k31s;
    }
    if (P.iH2shieldcustom > 0) k31s = c.fshield * k31s;
    RS(k31s) = k31s;
  }

  if (uvb_shielded(P)) {   // UVB self-shielding (:1492-1676)
    const double tfac = POW_POS(th.tgas / 1.0e4, 0.17);
    double nratio = (c.HI + c.HII);
    if (PC > 1) {
      nratio = nratio + c.HM + c.H2I + c.H2II;
      if (PC > 2) nratio = nratio + 0.5 * (c.DI + c.DII) + 2.0 * c.HDI / 3.0;
    }
    nratio = nratio * dom / (P.ssh_sig_hi * tfac * P.ssh_k24);
    const double fH = rahmati_f(P, nratio);
    nratio = 0.25 * (c.HeI + c.HeII + c.HeIII) * dom / (P.ssh_sig_hei * tfac * P.ssh_k26);
    const double fHe = rahmati_f(P, nratio);
    RS(k24s) = (P.k24 < GRB_TINY8) ? 0.0 : P.k24 * fH;
    RS(k29s) = (P.k29 < GRB_TINY8) ? 0.0 : P.k29 * fH;
    double k25s = P.k25, k26s = P.k26, k28s = P.k28, k30s = P.k30;
    if (P.iradshield == 2 || P.iradshield == 3) {
      k26s = (P.k26 < GRB_TINY8) ? 0.0 : P.k26 * fHe;
      k28s = (P.k28 < GRB_TINY8) ? 0.0 : P.k28 * fHe;
      k30s = (P.k30 < GRB_TINY8) ? 0.0 : P.k30 * fHe;
      k25s = (P.iradshield == 2) ? P.k25 : 0.0;
    }
    RS(k25s) = k25s; RS(k26s) = k26s; RS(k28s) = k28s; RS(k30s) = k30s;
  }
}

// ---------------------------------------------------------------------------------
// rate_timestep_g (solve_rate_cool_g.F:1743-1953): dedot, HIdot, and H2-formation
// heating added to edot.
// ---------------------------------------------------------------------------------
// k13 = the density-dependent rate of k13_martin when three_body_rate == 0, the tabulated column otherwise
template <int PC, bool FOLD, class R>
__device__ inline void rate_derivs(const GrParams &P, const Cell &c, const Thermo &th,
                                   const R &r, double k13, double &dedot, double &HIdot, double &edot)
{
  const bool anydust = (P.idust > 0) || (P.idustall > 0);
  const double de = c.de, HI = c.HI, HII = c.HII, HeI = c.HeI, HeII = c.HeII, HeIII = c.HeIII;
  if (PC == 1) {
    dedot = + RK(k1) * HI * de + RK(k3) * HeI * de / 4.0 + RK(k5) * HeII * de / 4.0
            - RK(k2) * HII * de - RK(k4) * HeII * de / 4.0 - RK(k6) * HeIII * de / 4.0
            + RK(k57) * HI * HI + RK(k58) * HI * HeI / 4.0
            + (Q_k24s * HI + Q_k25s * HeII / 4.0 + Q_k26s * HeI / 4.0);
    HIdot = - RK(k1) * HI * de + RK(k2) * HII * de - RK(k57) * HI * HI - RK(k58) * HI * HeI / 4.0 - Q_k24s * HI;
  } else {
    const double HM = c.HM, H2I = c.H2I, H2II = c.H2II;
    HIdot = - RK(k1) * de * HI - RK(k7) * de * HI - RK(k8) * HM * HI - RK(k9) * HII * HI
            - RK(k10) * H2II * HI / 2.0 - 2.0 * RK(k22) * (HI * HI) * HI
            + RK(k2) * HII * de + 2.0 * k13 * HI * H2I / 2.0
            + RK(k11) * HII * H2I / 2.0 + 2.0 * RK(k12) * de * H2I / 2.0
            + RK(k14) * HM * de + RK(k15) * HM * HI + 2.0 * RK(k16) * HM * HII
            + 2.0 * RK(k18) * H2II * de / 2.0 + RK(k19) * H2II * HM / 2.0
            - RK(k57) * HI * HI - RK(k58) * HI * HeI / 4.0 - Q_k24s * HI
            + 2.0 * Q_k31s * H2I / 2.0;
    if (anydust) HIdot = HIdot - 2.0 * Q_h2dust * th.rhoH;

    dedot = + RK(k1) * HI * de + RK(k3) * HeI * de / 4.0 + RK(k5) * HeII * de / 4.0
            + RK(k8) * HM * HI + RK(k15) * HM * HI + RK(k17) * HM * HII + RK(k14) * HM * de
            - RK(k2) * HII * de - RK(k4) * HeII * de / 4.0 - RK(k6) * HeIII * de / 4.0
            - RK(k7) * HI * de - RK(k18) * H2II * de / 2.0
            + RK(k57) * HI * HI + RK(k58) * HI * HeI / 4.0
            + (Q_k24s * HI + Q_k25s * HeII / 4.0 + Q_k26s * HeI / 4.0);

    // H2 formation heating, Omukai (2000) eq. 23
    double h2heatfac;
    {
      const double ncr_den = P.dom * (HI * RK(ncrd1) + H2I * 0.5 * RK(ncrd2));
      FmFast f{P.fm_tab, false};
      h2heatfac = f.div_s(1.0, 1.0 + f.div_s(RK(ncrn), ncr_den));
      if (f.bad) h2heatfac = ddiv_s(1.0, 1.0 + ddiv_s(RK(ncrn), ncr_den));
    }
    double H2delta = HI * (4.48 * RK(k22) * (HI * HI) - 4.48 * k13 * H2I / 2.0);
    if (H2delta > 0.0) H2delta = H2delta * h2heatfac;
    if (anydust) H2delta = H2delta + Q_h2dust * HI * th.rhoH * (0.2 + 4.2 * h2heatfac);
    edot = edot + P.chunit * H2delta * 1.0;
  }
  if (P.iradtrans == 1) {
    HIdot = HIdot - c.kphHI * HI;
    if (P.irt_honly == 0) dedot = dedot + c.kphHI * HI + c.kphHeI * HeI / 4.0 + c.kphHeII * HeII / 4.0;
    else dedot = dedot + c.kphHI * HI;
  }
}

// ---------------------------------------------------------------------------------
// step_rate_g (solve_rate_cool_g.F:1961-2413): one linearly-implicit Gauss-Seidel
// backward-Euler sweep.  Update order matters; electrons come from charge conservation
// using the OLD HM/H2II (:2379-2382).  Returns |dHI|/dt and |dde|/dt for the iter>50 rule.
// ---------------------------------------------------------------------------------
template <int PC, bool FOLD, class R>
__device__ inline void step_species(const GrParams &P, Cell &c, const Thermo &th, const R &r,
                                    double k13, double dt, double &HIdot_prev, double &dedot_prev)
{
  const bool anydust = (P.idust > 0) || (P.idustall > 0);
  const bool rt = (P.iradtrans == 1);
  const bool rt_all = rt && (P.irt_honly == 0);
  const bool rt_honly = rt && (P.irt_honly == 1);
  const double de = c.de, HI = c.HI, HII = c.HII, HeI = c.HeI, HeII = c.HeII, HeIII = c.HeIII;
  const double kphHI = rt ? c.kphHI : 0.0;
  const double kphHeI = rt_all ? c.kphHeI : 0.0;
  const double kphHeII = rt_all ? c.kphHeII : 0.0;
  double scoef, acoef;
  double HIp = 0, HIIp = 0, dep = 0, HMp = 0, H2Ip = 0, H2IIp = 0, DIp = 0, DIIp = 0, HDIp = 0;

  if (PC == 1) {
    scoef = RK(k2) * HII * de;
    acoef = RK(k1) * de + RK(k57) * HI + RK(k58) * HeI / 4.0 + Q_k24s;
    if (rt) acoef = acoef + kphHI;
    HIp = ddiv((scoef * dt + HI), (1.0 + acoef * dt));

    scoef = RK(k1) * HIp * de + RK(k57) * HIp * HIp + RK(k58) * HIp * HeI / 4.0 + Q_k24s * HIp;
    if (rt) scoef = scoef + kphHI * HIp;
    acoef = RK(k2) * de;
    HIIp = ddiv((scoef * dt + HII), (1.0 + acoef * dt));
    // (the reference also forms dep here and then discards it)
  }

  // helium, every network
  scoef = RK(k4) * HeII * de;
  acoef = RK(k3) * de + Q_k26s;
  if (rt_all) acoef = acoef + kphHeI;
  const double HeIp = ddiv((scoef * dt + HeI), (1.0 + acoef * dt));

  scoef = RK(k3) * HeIp * de + RK(k6) * HeIII * de + Q_k26s * HeIp;
  if (rt_all) scoef = scoef + kphHeI * HeIp;
  acoef = RK(k4) * de + RK(k5) * de + Q_k25s;
  if (rt_all) acoef = acoef + kphHeII;
  const double HeIIp = ddiv((scoef * dt + HeII), (1.0 + acoef * dt));

  scoef = RK(k5) * HeIIp * de + Q_k25s * HeIIp;
  if (rt_all) scoef = scoef + kphHeII * HeIIp;
  acoef = RK(k6) * de;
  const double HeIIIp = ddiv((scoef * dt + HeIII), (1.0 + acoef * dt));

  if (PC > 1) {
    const double HM = c.HM, H2I = c.H2I, H2II = c.H2II;
    scoef = RK(k2) * HII * de + 2.0 * k13 * HI * H2I / 2.0 + RK(k11) * HII * H2I / 2.0 +
            2.0 * RK(k12) * de * H2I / 2.0 + RK(k14) * HM * de + RK(k15) * HM * HI +
            2.0 * RK(k16) * HM * HII + 2.0 * RK(k18) * H2II * de / 2.0 + RK(k19) * H2II * HM / 2.0 +
            2.0 * Q_k31s * H2I / 2.0;
    acoef = RK(k1) * de + RK(k7) * de + RK(k8) * HM + RK(k9) * HII + RK(k10) * H2II / 2.0 +
            2.0 * RK(k22) * (HI * HI) + RK(k57) * HI + RK(k58) * HeI / 4.0 + Q_k24s;
    if (rt) acoef = acoef + kphHI;
    if (anydust) acoef = acoef + 2.0 * Q_h2dust * th.rhoH;
    HIp = ddiv((scoef * dt + HI), (1. + acoef * dt));

    scoef = RK(k1) * HI * de + RK(k10) * H2II * HI / 2.0 + RK(k57) * HI * HI + RK(k58) * HI * HeI / 4.0 +
            Q_k24s * HI;
    if (rt) scoef = scoef + kphHI * HI;
    acoef = RK(k2) * de + RK(k9) * HI + RK(k11) * H2I / 2.0 + RK(k16) * HM + RK(k17) * HM;
    HIIp = ddiv((scoef * dt + HII), (1.0 + acoef * dt));

    scoef = RK(k8) * HM * HI + RK(k15) * HM * HI + RK(k17) * HM * HII + RK(k57) * HI * HI +
            RK(k58) * HI * HeI / 4.0 + Q_k24s * HIp + Q_k25s * HeIIp / 4.0 + Q_k26s * HeIp / 4.0;
    if (rt_all) scoef = scoef + kphHI * HIp + kphHeI * HeIp / 4.0 + kphHeII * HeIIp / 4.0;
    if (rt_honly) scoef = scoef + kphHI * HIp;
    acoef = -(RK(k1) * HI - RK(k2) * HII + RK(k3) * HeI / 4.0 - RK(k6) * HeIII / 4.0 + RK(k5) * HeII / 4.0 -
              RK(k4) * HeII / 4.0 + RK(k14) * HM - RK(k7) * HI - RK(k18) * H2II / 2.0);
    dep = ddiv(scoef * dt + de, 1.0 + acoef * dt);

    scoef = 2.0 * (RK(k8) * HM * HI + RK(k10) * H2II * HI / 2.0 + RK(k19) * H2II * HM / 2.0 +
                   RK(k22) * HI * (HI * HI));
    acoef = (k13 * HI + RK(k11) * HII + RK(k12) * de) + Q_k29s + Q_k31s;
    if (anydust) scoef = scoef + 2.0 * Q_h2dust * HI * th.rhoH;
    H2Ip = ddiv((scoef * dt + H2I), (1.0 + acoef * dt));

    scoef = RK(k7) * HI * de;
    acoef = (RK(k8) + RK(k15)) * HI + (RK(k16) + RK(k17)) * HII + RK(k14) * de + RK(k19) * H2II / 2.0 + P.k27;
    HMp = ddiv((scoef * dt + HM), (1.0 + acoef * dt));

    // H2+ at equilibrium with the NEW abundances (:2293-2301)
    H2IIp = ddiv(2.0 * (RK(k9) * HIp * HIIp + RK(k11) * H2Ip / 2.0 * HIIp + RK(k17) * HMp * HIIp + Q_k29s * H2Ip),
                 RK(k10) * HIp + RK(k18) * dep + RK(k19) * HMp + (Q_k28s + Q_k30s));
  }

  if (PC > 2) {
    const double HM = c.HM, H2I = c.H2I, DI = c.DI, DII = c.DII, HDI = c.HDI;
    scoef = (RK(k2) * DII * de + RK(k51) * DII * HI + 2.0 * RK(k55) * HDI * HI * (1.0 / 3.0));
    acoef = RK(k1) * de + RK(k50) * HII + RK(k54) * H2I / 2.0 + RK(k56) * HM + Q_k24s;
    if (rt) acoef = acoef + kphHI;
    DIp = ddiv((scoef * dt + DI), (1.0 + acoef * dt));

    scoef = (RK(k1) * DI * de + RK(k50) * HII * DI + 2.0 * RK(k53) * HII * HDI * (1.0 / 3.0)) + Q_k24s * DI;
    if (rt) scoef = scoef + kphHI * DI;
    acoef = RK(k2) * de + RK(k51) * HI + RK(k52) * H2I / 2.0;
    DIIp = ddiv((scoef * dt + DII), (1.0 + acoef * dt));

    scoef = 3.0 * (RK(k52) * DII * H2I / 2.0 / 2.0 + RK(k54) * DI * H2I / 2.0 / 2.0 +
                   2.0 * RK(k56) * DI * HM / 2.0);
    acoef = RK(k53) * HII + RK(k55) * HI;
    HDIp = ddiv((scoef * dt + HDI), (1.0 + acoef * dt));
  }

  // write back with floors (:2364-2397)
  const double rdt = drcp(dmax(dt, GRB_TINY8));
  HIdot_prev = fabs(c.HI - HIp) * rdt;
  c.HI = dmax(HIp, GRB_TINY);
  c.HII = dmax(HIIp, GRB_TINY);
  c.HeI = dmax(HeIp, GRB_TINY);
  c.HeII = dmax(HeIIp, GRB_TINY);
  c.HeIII = dmax(HeIIIp, 1e-5 * GRB_TINY);
  const double de_old = c.de;
  c.de = c.HII + c.HeII / 4.0 + c.HeIII / 2.0;
  if (PC > 1) c.de = c.de - c.HM + c.H2II / 2.0;     // old HM, H2II
  dedot_prev = fabs(c.de - de_old) * rdt;
  if (PC > 1) {
    c.HM = dmax(HMp, GRB_TINY);
    c.H2I = dmax(H2Ip, GRB_TINY);
    c.H2II = dmax(H2IIp, GRB_TINY);
  }
  if (PC > 2) {
    c.DI = dmax(DIp, GRB_TINY);
    c.DII = dmax(DIIp, GRB_TINY);
    c.HDI = dmax(HDIp, GRB_TINY);
  }
}

// make_consistent_g (solve_rate_cool_g.F:2419-2535) for one cell
template <int PC>
__device__ inline void make_consistent(const GrParams &P, Cell &c)
{
  const double metalfree = (P.imetal == 1) ? c.d - c.metal : c.d;
  c.HI = fabs(c.HI); c.HII = fabs(c.HII); c.HeI = fabs(c.HeI); c.HeII = fabs(c.HeII); c.HeIII = fabs(c.HeIII);
  double totalH = c.HI + c.HII;
  const double totalHe = c.HeI + c.HeII + c.HeIII;
  if (PC > 1) {
    c.HM = fabs(c.HM); c.H2II = fabs(c.H2II); c.H2I = fabs(c.H2I);
    totalH = totalH + c.HM + c.H2I + c.H2II;
  }
  const double correctH = ddiv(P.fh * metalfree, totalH);
  c.HI = c.HI * correctH;
  c.HII = c.HII * correctH;
  const double correctHe = ddiv((1.0 - P.fh) * metalfree, totalHe);
  c.HeI = c.HeI * correctHe; c.HeII = c.HeII * correctHe; c.HeIII = c.HeIII * correctHe;
  if (PC > 1) { c.HM = c.HM * correctH; c.H2II = c.H2II * correctH; c.H2I = c.H2I * correctH; }
  if (PC > 2) {
    c.DI = fabs(c.DI); c.DII = fabs(c.DII); c.HDI = fabs(c.HDI);
    const double totalD = c.DI + c.DII + 2.0 / 3.0 * c.HDI;
    const double correctD = ddiv(P.fh * P.dtoh * metalfree, totalD);
    c.DI = c.DI * correctD; c.DII = c.DII * correctD; c.HDI = c.HDI * correctD;
  }
  c.de = c.HII + c.HeII / 4.0 + c.HeIII / 2.0;
  if (PC > 1) c.de = c.de - c.HM + c.H2II / 2.0;
}

}  // namespace grb
#endif
