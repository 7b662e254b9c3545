// grackle-b200: rate lookup, rate-limited time step and the backward-Euler species
// update for one cell (continuation of chem.cuh).
#ifndef GRB200_NETWORK_CUH
#define GRB200_NETWORK_CUH

#include "chem.cuh"

namespace grb {

struct Rates {
  double k1, k2, k3, k4, k5, k6, k57, k58;
  double k7, k8, k9, k10, k11, k12, k13, k14, k15, k16, k17, k18, k19, k22;
  double ncrn, ncrd1, ncrd2;
  double k50, k51, k52, k53, k54, k55, k56;
  double h2dust;
  double k24s, k25s, k26s, k28s, k29s, k30s, k31s;   // (self-)shielded photo rates
  // log-T bracket of this lookup, needed by the H-equilibrium dt limiter
  int idx;
  double t1, t2;
};

// neighbour densities for the Sobolev-like H2 shielding length (H2_self_shielding == 1)
struct Neigh { double d[6]; };

// ---------------------------------------------------------------------------------
// lookup_cool_rates1d_g (solve_rate_cool_g.F:1079-1737)
// ---------------------------------------------------------------------------------
template <int PC>
__device__ inline void lookup_rates(const GrParams &P, const Cell &c, const Thermo &th,
                                    const Neigh &nb, Rates &r)
{
  const bool anydust = (P.idust > 0) || (P.idustall > 0);
  const double dom = P.dom;
  double logtem = log(th.tgas);
  logtem = dmax(logtem, P.logtem0);
  logtem = dmin(logtem, P.logtem9);
  const TIndex ti = t_index(P, logtem);
  r.idx = ti.idx; r.t1 = ti.t1; r.t2 = ti.t2;
  const double tdef = ti.tdef;
  const double *row = P.chem + (size_t)(ti.idx - 1) * NCHEM;
#define K_LERP(col) lerp_row(row, NCHEM, col, tdef)

  r.k1 = K_LERP(C_K1); r.k2 = K_LERP(C_K2); r.k3 = K_LERP(C_K3); r.k4 = K_LERP(C_K4);
  r.k5 = K_LERP(C_K5); r.k6 = K_LERP(C_K6); r.k57 = K_LERP(C_K57); r.k58 = K_LERP(C_K58);

  if (PC > 1) {
    r.k7 = K_LERP(C_K7); r.k8 = K_LERP(C_K8); r.k9 = K_LERP(C_K9); r.k10 = K_LERP(C_K10);
    r.k11 = K_LERP(C_K11); r.k12 = K_LERP(C_K12); r.k13 = K_LERP(C_K13); r.k14 = K_LERP(C_K14);
    r.k15 = K_LERP(C_K15); r.k16 = K_LERP(C_K16); r.k17 = K_LERP(C_K17); r.k18 = K_LERP(C_K18);
    r.k19 = K_LERP(C_K19); r.k22 = K_LERP(C_K22);
    r.ncrn = K_LERP(C_NCRN); r.ncrd1 = K_LERP(C_NCRD1); r.ncrd2 = K_LERP(C_NCRD2);
  }
  if (PC > 2) {
    r.k50 = K_LERP(C_K50); r.k51 = K_LERP(C_K51); r.k52 = K_LERP(C_K52); r.k53 = K_LERP(C_K53);
    r.k54 = K_LERP(C_K54); r.k55 = K_LERP(C_K55); r.k56 = K_LERP(C_K56);
  }

  // H2 formation on grains: bilinear in (ln T, ln T_dust) (:1327-1378)
  r.h2dust = 0.0;
  if (anydust) {
    if (th.tdust > P.dtemend) {
      r.h2dust = GRB_TINY8;
    } else {
      double dl = log(th.tdust);
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
      r.h2dust = dusti1 + (dusti2 - dusti1) * d_tdef;
      r.h2dust = r.h2dust * th.dust2gas;
    }
  }

  r.k24s = P.k24; r.k25s = P.k25; r.k26s = P.k26; r.k28s = P.k28; r.k29s = P.k29; r.k30s = P.k30;
  r.k31s = 0.0;

  if (PC > 1) {     // H2 self-shielding, Wolcott-Green & Haiman (2019) (:1397-1490)
    r.k31s = (P.iradtrans == 0) ? P.k31 : P.k31 + c.kdissH2I;
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
      const double aWG = (0.8711 * log10(tuse) - 1.928) * exp(-0.2856 * log10(nuse)) +
                         (-0.9639 * log10(tuse) + 3.892);
      const double x = 2.0e-15 * N_H2;
      const double b_doppler = 1e-5 * sqrt(2.0 * GRB_KBOLTZ * th.tgas / GRB_MASS_H);
      double f_shield = 0.965 / pow(1.0 + x / b_doppler, aWG) +
                        0.035 * exp(-8.5e-4 * sqrt(1.0 + x)) / sqrt(1.0 + x);
      f_shield = dmin(f_shield, 1.0);
      r.k31s = f_shield * r.k31s;
    }
    if (P.iH2shieldcustom > 0) r.k31s = c.fshield * r.k31s;
  }

  if (P.iradshield > 0) {   // UVB self-shielding (:1492-1676)
    const double tfac = pow(th.tgas / 1.0e4, 0.17);
    double nratio = (c.HI + c.HII);
    if (PC > 1) {
      nratio = nratio + c.HM + c.H2I + c.H2II;
      if (PC > 2) nratio = nratio + 0.5 * (c.DI + c.DII) + 2.0 * c.HDI / 3.0;
    }
    nratio = nratio * dom / (P.ssh_sig_hi * tfac * P.ssh_k24);
    const double fH = rahmati_f(nratio);
    nratio = 0.25 * (c.HeI + c.HeII + c.HeIII) * dom / (P.ssh_sig_hei * tfac * P.ssh_k26);
    const double fHe = rahmati_f(nratio);
    r.k24s = (P.k24 < GRB_TINY8) ? 0.0 : r.k24s * fH;
    r.k29s = (P.k29 < GRB_TINY8) ? 0.0 : r.k29s * fH;
    if (P.iradshield == 1) {
      r.k25s = P.k25; r.k26s = P.k26;
    } else if (P.iradshield == 2 || P.iradshield == 3) {
      r.k26s = (P.k26 < GRB_TINY8) ? 0.0 : r.k26s * fHe;
      r.k28s = (P.k28 < GRB_TINY8) ? 0.0 : r.k28s * fHe;
      r.k30s = (P.k30 < GRB_TINY8) ? 0.0 : r.k30s * fHe;
      r.k25s = (P.iradshield == 2) ? P.k25 : 0.0;
    }
  }

  // density-dependent collisional dissociation of H2 (Martin+96) (:1709-1733)
  if (PC > 1 && P.ithreebody == 0) {
    const double nh = dmin(c.HI * dom, 1.0e9);
    r.k13 = GRB_TINY8;
    if (th.tgas >= 500.0 && th.tgas < 1.0e6) {
      double dd[14];
#pragma unroll
      for (int n = 0; n < 14; n++) dd[n] = K_LERP(C_K13DD0 + n);
      double cid = dd[0] - dd[1] / (1.0 + pow(nh / dd[4], dd[6])) + dd[2] -
                   dd[3] / (1.0 + pow(nh / dd[5], dd[6]));
      cid = dmax(pow(10.0, cid), GRB_TINY8);
      double dtn = dd[7] - dd[8] / (1.0 + pow(nh / dd[11], dd[13])) + dd[9] -
                   dd[10] / (1.0 + pow(nh / dd[12], dd[13]));
      dtn = dmax(pow(10.0, dtn), GRB_TINY8);
      r.k13 = dtn + cid;
    }
  }
#undef K_LERP
}

// ---------------------------------------------------------------------------------
// rate_timestep_g (solve_rate_cool_g.F:1743-1953): dedot, HIdot, and H2-formation
// heating added to edot.
// ---------------------------------------------------------------------------------
template <int PC>
__device__ inline void rate_derivs(const GrParams &P, const Cell &c, const Thermo &th,
                                   const Rates &r, double &dedot, double &HIdot, double &edot)
{
  const bool anydust = (P.idust > 0) || (P.idustall > 0);
  const double de = c.de, HI = c.HI, HII = c.HII, HeI = c.HeI, HeII = c.HeII, HeIII = c.HeIII;
  if (PC == 1) {
    dedot = + r.k1 * HI * de + r.k3 * HeI * de / 4.0 + r.k5 * HeII * de / 4.0
            - r.k2 * HII * de - r.k4 * HeII * de / 4.0 - r.k6 * HeIII * de / 4.0
            + r.k57 * HI * HI + r.k58 * HI * HeI / 4.0
            + (r.k24s * HI + r.k25s * HeII / 4.0 + r.k26s * HeI / 4.0);
    HIdot = - r.k1 * HI * de + r.k2 * HII * de - r.k57 * HI * HI - r.k58 * HI * HeI / 4.0 - r.k24s * HI;
  } else {
    const double HM = c.HM, H2I = c.H2I, H2II = c.H2II;
    HIdot = - r.k1 * de * HI - r.k7 * de * HI - r.k8 * HM * HI - r.k9 * HII * HI
            - r.k10 * H2II * HI / 2.0 - 2.0 * r.k22 * (HI * HI) * HI
            + r.k2 * HII * de + 2.0 * r.k13 * HI * H2I / 2.0
            + r.k11 * HII * H2I / 2.0 + 2.0 * r.k12 * de * H2I / 2.0
            + r.k14 * HM * de + r.k15 * HM * HI + 2.0 * r.k16 * HM * HII
            + 2.0 * r.k18 * H2II * de / 2.0 + r.k19 * H2II * HM / 2.0
            - r.k57 * HI * HI - r.k58 * HI * HeI / 4.0 - r.k24s * HI
            + 2.0 * r.k31s * H2I / 2.0;
    if (anydust) HIdot = HIdot - 2.0 * r.h2dust * th.rhoH;

    dedot = + r.k1 * HI * de + r.k3 * HeI * de / 4.0 + r.k5 * HeII * de / 4.0
            + r.k8 * HM * HI + r.k15 * HM * HI + r.k17 * HM * HII + r.k14 * HM * de
            - r.k2 * HII * de - r.k4 * HeII * de / 4.0 - r.k6 * HeIII * de / 4.0
            - r.k7 * HI * de - r.k18 * H2II * de / 2.0
            + r.k57 * HI * HI + r.k58 * HI * HeI / 4.0
            + (r.k24s * HI + r.k25s * HeII / 4.0 + r.k26s * HeI / 4.0);

    // H2 formation heating, Omukai (2000) eq. 23
    const double h2heatfac = 1.0 / (1.0 + (r.ncrn / (P.dom * (HI * r.ncrd1 + H2I * 0.5 * r.ncrd2))));
    double H2delta = HI * (4.48 * r.k22 * (HI * HI) - 4.48 * r.k13 * H2I / 2.0);
    if (H2delta > 0.0) H2delta = H2delta * h2heatfac;
    if (anydust) H2delta = H2delta + r.h2dust * HI * th.rhoH * (0.2 + 4.2 * h2heatfac);
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
template <int PC>
__device__ inline void step_species(const GrParams &P, Cell &c, const Thermo &th, const Rates &r,
                                    double dt, double &HIdot_prev, double &dedot_prev)
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
    scoef = r.k2 * HII * de;
    acoef = r.k1 * de + r.k57 * HI + r.k58 * HeI / 4.0 + r.k24s;
    if (rt) acoef = acoef + kphHI;
    HIp = (scoef * dt + HI) / (1.0 + acoef * dt);

    scoef = r.k1 * HIp * de + r.k57 * HIp * HIp + r.k58 * HIp * HeI / 4.0 + r.k24s * HIp;
    if (rt) scoef = scoef + kphHI * HIp;
    acoef = r.k2 * de;
    HIIp = (scoef * dt + HII) / (1.0 + acoef * dt);
    // (the reference also forms dep here and then discards it)
  }

  // helium, every network
  scoef = r.k4 * HeII * de;
  acoef = r.k3 * de + r.k26s;
  if (rt_all) acoef = acoef + kphHeI;
  const double HeIp = (scoef * dt + HeI) / (1.0 + acoef * dt);

  scoef = r.k3 * HeIp * de + r.k6 * HeIII * de + r.k26s * HeIp;
  if (rt_all) scoef = scoef + kphHeI * HeIp;
  acoef = r.k4 * de + r.k5 * de + r.k25s;
  if (rt_all) acoef = acoef + kphHeII;
  const double HeIIp = (scoef * dt + HeII) / (1.0 + acoef * dt);

  scoef = r.k5 * HeIIp * de + r.k25s * HeIIp;
  if (rt_all) scoef = scoef + kphHeII * HeIIp;
  acoef = r.k6 * de;
  const double HeIIIp = (scoef * dt + HeIII) / (1.0 + acoef * dt);

  if (PC > 1) {
    const double HM = c.HM, H2I = c.H2I, H2II = c.H2II;
    scoef = r.k2 * HII * de + 2.0 * r.k13 * HI * H2I / 2.0 + r.k11 * HII * H2I / 2.0 +
            2.0 * r.k12 * de * H2I / 2.0 + r.k14 * HM * de + r.k15 * HM * HI +
            2.0 * r.k16 * HM * HII + 2.0 * r.k18 * H2II * de / 2.0 + r.k19 * H2II * HM / 2.0 +
            2.0 * r.k31s * H2I / 2.0;
    acoef = r.k1 * de + r.k7 * de + r.k8 * HM + r.k9 * HII + r.k10 * H2II / 2.0 +
            2.0 * r.k22 * (HI * HI) + r.k57 * HI + r.k58 * HeI / 4.0 + r.k24s;
    if (rt) acoef = acoef + kphHI;
    if (anydust) acoef = acoef + 2.0 * r.h2dust * th.rhoH;
    HIp = (scoef * dt + HI) / (1. + acoef * dt);

    scoef = r.k1 * HI * de + r.k10 * H2II * HI / 2.0 + r.k57 * HI * HI + r.k58 * HI * HeI / 4.0 +
            r.k24s * HI;
    if (rt) scoef = scoef + kphHI * HI;
    acoef = r.k2 * de + r.k9 * HI + r.k11 * H2I / 2.0 + r.k16 * HM + r.k17 * HM;
    HIIp = (scoef * dt + HII) / (1.0 + acoef * dt);

    scoef = r.k8 * HM * HI + r.k15 * HM * HI + r.k17 * HM * HII + r.k57 * HI * HI +
            r.k58 * HI * HeI / 4.0 + r.k24s * HIp + r.k25s * HeIIp / 4.0 + r.k26s * HeIp / 4.0;
    if (rt_all) scoef = scoef + kphHI * HIp + kphHeI * HeIp / 4.0 + kphHeII * HeIIp / 4.0;
    if (rt_honly) scoef = scoef + kphHI * HIp;
    acoef = -(r.k1 * HI - r.k2 * HII + r.k3 * HeI / 4.0 - r.k6 * HeIII / 4.0 + r.k5 * HeII / 4.0 -
              r.k4 * HeII / 4.0 + r.k14 * HM - r.k7 * HI - r.k18 * H2II / 2.0);
    dep = (scoef * dt + de) / (1.0 + acoef * dt);

    scoef = 2.0 * (r.k8 * HM * HI + r.k10 * H2II * HI / 2.0 + r.k19 * H2II * HM / 2.0 +
                   r.k22 * HI * (HI * HI));
    acoef = (r.k13 * HI + r.k11 * HII + r.k12 * de) + r.k29s + r.k31s;
    if (anydust) scoef = scoef + 2.0 * r.h2dust * HI * th.rhoH;
    H2Ip = (scoef * dt + H2I) / (1.0 + acoef * dt);

    scoef = r.k7 * HI * de;
    acoef = (r.k8 + r.k15) * HI + (r.k16 + r.k17) * HII + r.k14 * de + r.k19 * H2II / 2.0 + P.k27;
    HMp = (scoef * dt + HM) / (1.0 + acoef * dt);

    // H2+ at equilibrium with the NEW abundances (:2293-2301)
    H2IIp = 2.0 * (r.k9 * HIp * HIIp + r.k11 * H2Ip / 2.0 * HIIp + r.k17 * HMp * HIIp + r.k29s * H2Ip) /
            (r.k10 * HIp + r.k18 * dep + r.k19 * HMp + (r.k28s + r.k30s));
  }

  if (PC > 2) {
    const double HM = c.HM, H2I = c.H2I, DI = c.DI, DII = c.DII, HDI = c.HDI;
    scoef = (r.k2 * DII * de + r.k51 * DII * HI + 2.0 * r.k55 * HDI * HI / 3.0);
    acoef = r.k1 * de + r.k50 * HII + r.k54 * H2I / 2.0 + r.k56 * HM + r.k24s;
    if (rt) acoef = acoef + kphHI;
    DIp = (scoef * dt + DI) / (1.0 + acoef * dt);

    scoef = (r.k1 * DI * de + r.k50 * HII * DI + 2.0 * r.k53 * HII * HDI / 3.0) + r.k24s * DI;
    if (rt) scoef = scoef + kphHI * DI;
    acoef = r.k2 * de + r.k51 * HI + r.k52 * H2I / 2.0;
    DIIp = (scoef * dt + DII) / (1.0 + acoef * dt);

    scoef = 3.0 * (r.k52 * DII * H2I / 2.0 / 2.0 + r.k54 * DI * H2I / 2.0 / 2.0 +
                   2.0 * r.k56 * DI * HM / 2.0);
    acoef = r.k53 * HII + r.k55 * HI;
    HDIp = (scoef * dt + HDI) / (1.0 + acoef * dt);
  }

  // write back with floors (:2364-2397)
  HIdot_prev = fabs(c.HI - HIp) / dmax(dt, GRB_TINY8);
  c.HI = dmax(HIp, GRB_TINY);
  c.HII = dmax(HIIp, GRB_TINY);
  c.HeI = dmax(HeIp, GRB_TINY);
  c.HeII = dmax(HeIIp, GRB_TINY);
  c.HeIII = dmax(HeIIIp, 1e-5 * GRB_TINY);
  const double de_old = c.de;
  c.de = c.HII + c.HeII / 4.0 + c.HeIII / 2.0;
  if (PC > 1) c.de = c.de - c.HM + c.H2II / 2.0;     // old HM, H2II
  dedot_prev = fabs(c.de - de_old) / dmax(dt, GRB_TINY8);
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
  const double correctH = P.fh * metalfree / totalH;
  c.HI = c.HI * correctH;
  c.HII = c.HII * correctH;
  const double correctHe = (1.0 - P.fh) * metalfree / totalHe;
  c.HeI = c.HeI * correctHe; c.HeII = c.HeII * correctHe; c.HeIII = c.HeIII * correctHe;
  if (PC > 1) { c.HM = c.HM * correctH; c.H2II = c.H2II * correctH; c.H2I = c.H2I * correctH; }
  if (PC > 2) {
    c.DI = fabs(c.DI); c.DII = fabs(c.DII); c.HDI = fabs(c.HDI);
    const double totalD = c.DI + c.DII + 2.0 / 3.0 * c.HDI;
    const double correctD = P.fh * P.dtoh * metalfree / totalD;
    c.DI = c.DI * correctD; c.DII = c.DII * correctD; c.HDI = c.HDI * correctD;
  }
  c.de = c.HII + c.HeII / 4.0 + c.HeIII / 2.0;
  if (PC > 1) c.de = c.de - c.HM + c.H2II / 2.0;
}

}  // namespace grb
#endif
