// grackle-b200: the per-cell cooling/heating rate and the chemistry network update.
// (continuation of physics.cuh; same conventions)
#ifndef GRB200_CHEM_CUH
#define GRB200_CHEM_CUH

#include "physics.cuh"

namespace grb {

// One cell, proper (not comoving) densities, all in registers.
struct Cell {
  double d, e, de, HI, HII, HeI, HeII, HeIII, HM, H2I, H2II, DI, DII, HDI, metal, dust;
  // optional per-cell source terms (0 when the physics is off)
  double vheat, mheat, kphHI, kphHeI, kphHeII, kdissH2I, photogamma, xH2shield, isrf, fshield;
};

// what cool_rate leaves behind for the rest of the subcycle (the reference's row temporaries)
struct Thermo {
  double edot, tgas, mmw, p2d, tdust, metallicity, dust2gas, rhoH, myde;
  double logtgas;     // ln(tgas), formed once here and reused by the rate lookup
};

// One row of a packed table holds (value, difference to the next bin) pairs: column `col`
// interpolated at weight t is a + d * t - the reference's a + (b - a) * t
// (cool1d_multi_g.F:387-410, solve_rate_cool_g.F:1225-1240) with b - a formed on the host.
__device__ __forceinline__ double lerp_row(const double *row, int col, double t)
{
  const double2 p = __ldg(reinterpret_cast<const double2 *>(row) + col);
  return p.x + p.y * t;
}

// N consecutive columns starting at C0: all loads are issued before the first use
template <int C0, int N>
__device__ __forceinline__ void lerp_cols(const double *row, double t, double (&out)[N])
{
  double2 p[N];
  const double2 *q = reinterpret_cast<const double2 *>(row) + C0;
#pragma unroll
  for (int i = 0; i < N; i++) p[i] = __ldg(q + i);
#pragma unroll
  for (int i = 0; i < N; i++) out[i] = p[i].x + p[i].y * t;
}

// Rahmati+13 eq. 13/14 shielding fraction (cool1d_multi_g.F:788-792, solve_rate_cool_g.F:1518-1520)
__device__ __forceinline__ double rahmati_f(const GrParams &P, double nratio)
{
  return 0.98 * POW_POS(1.0 + POW_POS(nratio, 1.64), -2.28) + 0.02 * POW_POS(1.0 + nratio, -0.84);
}

struct TIndex { int idx; double t1, t2, tdef; };   // idx is 1-based like the Fortran

__device__ __forceinline__ TIndex t_index(const GrParams &P, double logtem)
{
  TIndex r;
  // (x - x0) / dlogtem and (x - t1) / (t2 - t1) as multiplications by 1/dlogtem: t2 - t1 is
  // dlogtem to within one rounding, and the interpolant is continuous across a bin edge
  r.idx = min(P.nratec - 1, max(1, (int)((logtem - P.logtem0) * P.inv_dlogtem) + 1));
  r.t1 = (P.logtem0 + (r.idx - 1) * P.dlogtem);
  r.t2 = (P.logtem0 + (r.idx) * P.dlogtem);
  r.tdef = (logtem - r.t1) * P.inv_dlogtem;
  return r;
}

// ---------------------------------------------------------------------------------
// cool1d_multi_g (cool1d_multi_g.F:6-1131) for one cell, in two parts so that the solve
// kernel can fetch the cooling-table rows between them:
//   thermo_head : temperature, mean molecular weight, ... (:212-357); returns the clamped
//                 ln of the temperature the cooling tables are read at
//   cool_sums   : every cooling / heating term (:360-1131).  CV supplies the interpolated
//                 value of cooling-table column `col` (registers for the one-shot
//                 calculate_* kernels, shared-memory slots in the solve kernel).
// `first` = (iter == 1).  tgasold is the per-cell carry between subcycles.
// ---------------------------------------------------------------------------------
// SET_Z = false: the caller has already put the cell's metallicity into th (the solve kernel
// forms it once per cell, not once per subcycle: metal and total density do not change)
template <int PC, bool SET_Z = true>
__device__ inline double thermo_head(const GrParams &P, const Cell &c, bool first, double &tgasold,
                                     Thermo &th, bool lookup_follows)
{
  th.p2d = (P.gamma - 1.0) * c.d * c.e;
  th.tdust = 0.0; th.dust2gas = 0.0;
  if (SET_Z) th.metallicity = 0.0;

  if (PC == 0) {
    th.rhoH = (P.imetal == 1) ? P.fh * (c.d - c.metal) : P.fh * c.d;
    temperature_table(P, c.d, c.e, c.metal, th.rhoH, th.tgas, th.mmw);
    th.myde = 0.0;
  } else {
    double mmw = (c.HeI + c.HeII + c.HeIII) / 4.0 + c.HI + c.HII + c.de;
    th.rhoH = c.HI + c.HII;
    th.myde = c.de;
    if (PC > 1) {
      mmw = mmw + c.HM + (c.H2I + c.H2II) / 2.0;
      th.rhoH = th.rhoH + c.H2I + c.H2II;
    }
    if (P.imetal == 1) mmw = mmw + c.metal / GRB_MU_METAL;
    const double rmmw = drcp(mmw);
    th.tgas = dmax(th.p2d * P.utem * rmmw, P.temstart);
    th.mmw = c.d * rmmw;
    if (PC > 1) {    // H2 vibrational gamma correction (:294-318)
      const double nH2 = 0.5 * (c.H2I + c.H2II);
      const double nother = (c.HeI + c.HeII + c.HeIII) / 4.0 + c.HI + c.HII + c.de;
      double gamma2;
      if (nH2 > 1.0e-3 * nother) {
        const double x = ddiv(6100.0, th.tgas);
        if (x > 10.0) {
          gamma2 = 0.5 * 5.0;
        } else {
          const double ex = DEXP(x);
          gamma2 = 0.5 * (5.0 + ddiv(2.0 * (x * x) * ex, (ex - 1) * (ex - 1)));
        }
      } else {
        gamma2 = 2.5;
      }
      gamma2 = 1.0 + ddiv(nH2 + nother, nH2 * gamma2 + nother * P.inv_gm1);
      th.tgas = th.tgas * (gamma2 - 1.0) * P.inv_gm1;
    }
  }

  if (SET_Z && P.imetal == 1) th.metallicity = ddiv(c.metal, c.d) * P.inv_zsolar;
  if (first) tgasold = th.tgas;

  // ln(tgas) for the rate lookup that follows in the solve (lookup_cool_rates1d_g :1211) and the
  // logarithm the cooling tables are read at: two independent evaluations, one basic block
  double logtem;
  {
    const double tmean = 0.5 * (th.tgas + tgasold);
    FmFast f{P.fm_tab, false};
    th.logtgas = (PC > 0 && lookup_follows) ? f.log(th.tgas) : 0.0;
    logtem = f.log(tmean);
    if (f.bad) {
      th.logtgas = (PC > 0 && lookup_follows) ? DLOG(th.tgas) : 0.0;
      logtem = DLOG(tmean);
    }
  }
  logtem = dmax(logtem, P.logtem0);
  logtem = dmin(logtem, P.logtem9);
  return logtem;
}

// cooling-table values of one cell held in registers (one-shot kernels)
struct CoolRegs {
  double v[NCOOL];
  __device__ __forceinline__ double operator()(int col) const { return v[col]; }
};

template <int PC, class CV>
__device__ inline void cool_sums(const GrParams &P, const Cell &c, double logtem, const CV &cv,
                                 double &tgasold, Thermo &th)
{
  const bool anydust = (P.idust > 0) || (P.idustall > 0) || (P.idustrec > 0);
  const double dom = P.dom, dom_inv = P.dom_inv;
  double edot = 0.0;
  const double mynh = th.rhoH * dom;

  if (PC > 0) {     // six-species atomic cooling (:417-446)
    const double ceHI = cv(L_CEHI), ceHeI = cv(L_CEHEI), ceHeII = cv(L_CEHEII);
    const double ciHI = cv(L_CIHI), ciHeI = cv(L_CIHEI), ciHeIS = cv(L_CIHEIS);
    const double ciHeII = cv(L_CIHEII), reHII = cv(L_REHII);
    const double reHeII1 = cv(L_REHEII1), reHeII2 = cv(L_REHEII2);
    const double reHeIII = cv(L_REHEIII), brem = cv(L_BREM);
    const double de = c.de;
    edot = (- ceHI * c.HI * de
            - ceHeI * c.HeII * (de * de) * dom / 4.0
            - ceHeII * c.HeII * de / 4.0
            - ciHI * c.HI * de
            - ciHeI * c.HeI * de / 4.0
            - ciHeII * c.HeII * de / 4.0
            - ciHeIS * c.HeII * (de * de) * dom / 4.0
            - reHII * c.HII * de
            - reHeII1 * c.HeII * de / 4.0
            - reHeII2 * c.HeII * de / 4.0
            - reHeIII * c.HeIII * de / 4.0
            - brem * (c.HII + c.HeII / 4.0 + c.HeIII) * de);
  }

  if (PC > 1) {     // H2 cooling, Glover & Abel (2008) (:472-526), CIE (:630-649)
    const double gaHI = cv(L_GAHI), gaH2 = cv(L_GAH2), gaHe = cv(L_GAHE), gaHp = cv(L_GAHP);
    const double gael = cv(L_GAEL), h2lte = cv(L_H2LTE);
    double fudge = 1.0;
    if (P.ih2optical == 1) {
      fudge = POW_POS(0.76 * c.d * dom / 8.e9, -0.45);
      fudge = dmin(fudge, 1.0);
    }
    const double galdl = gaHI * c.HI + gaH2 * c.H2I / 2.0 + gaHe * c.HeI / 4.0 + gaHp * c.HII + gael * c.de;
    const double gphdl1 = h2lte * dom_inv;
    {
      const double num = (double)P.ih2co * fudge * c.H2I * h2lte;
      FmFast f{P.fm_tab, false};
      double q = f.div_s(num, 1.0 + f.div_s(gphdl1, galdl));
      if (f.bad) q = ddiv_s(num, 1.0 + ddiv_s(gphdl1, galdl));
      edot = edot - q * (0.5 * dom_inv);
    }
    if (P.iciecool == 1 && c.d * dom > 1e10) {
      const double cieco = cv(L_CIECO);
      double tau = POW_POS((c.d / 2e16) * dom, 2.8);
      tau = dmax(tau, 1.e-5);
      double ciefudge = dmin((1.0 - DEXP(-tau)) / tau, 1.0);
      tau = POW_POS((c.d / 2.e18) * dom, 8.0);
      tau = dmax(tau, 1.e-5);
      ciefudge = ciefudge * dmin((1.0 - DEXP(-tau)) / tau, 1.0);
      edot = ciefudge * (edot - c.H2I * (c.d * cieco));
    }
  }

  if (PC > 2) {     // HD cooling (:655-686)
    double hdlte, hdlow;
    if (th.tgas > P.comp2) { hdlte = cv(L_HDLTE); hdlow = cv(L_HDLOW); }
    else { hdlte = GRB_TINY; hdlow = GRB_TINY; }
    const double hdlte1 = ddiv(hdlte, c.HI * dom);
    const double hdlow1 = dmax(hdlow, GRB_TINY);
    edot = edot - c.HDI * ddiv(hdlte, 1.0 + ddiv(hdlte1, hdlow1)) * ((1.0 / 3.0) * dom_inv);
  }

  if (anydust || (P.igammah > 0))
    th.dust2gas = (P.idustfield > 0) ? c.dust / c.d : P.fgr * th.metallicity;
  double myisrf = 0.0;
  if (anydust || (P.igammah > 1)) myisrf = (P.iisrffield > 0) ? c.isrf : P.isrf;

  if (anydust) {    // gas-grain heat transfer and T_dust (:726-753)
    const double gasgr = cv(L_GASGR);
    const double gasgr_tdust = P.fgr * gasgr * P.coolunit / GRB_MASS_H;
    th.tdust = dust_temperature(P, th.tgas, mynh, gasgr_tdust, P.gamma_isrfa, myisrf, P.comp2);
    edot = edot - gasgr * (th.tgas - th.tdust) * th.dust2gas * th.rhoH * th.rhoH;
  }

  if (PC > 0) {     // UVB photo-heating with optional self-shielding (:758-913)
    double fH = 1.0, fHe = 1.0;
    if (P.iradshield >= 1 && !(P.k24 < GRB_TINY8)) {
      const double nSSh = P.ssh_sig_hi * POW_POS(th.tgas / 1.0e4, 0.17) * P.ssh_k24;
      fH = rahmati_f(P, (c.HI + c.HII) * dom / nSSh);
    }
    if (P.iradshield >= 2 && !(P.k26 < GRB_TINY8)) {
      const double nSSh_he = P.ssh_sig_hei * POW_POS(th.tgas / 1.0e4, 0.17) * P.ssh_k26;
      fHe = rahmati_f(P, 0.25 * (c.HeI + c.HeII + c.HeIII) * dom / nSSh_he);
    }
    const double ip = (double)P.ipiht;
    if (P.iradshield == 0)
      edot = edot + ip * (P.piHI * c.HI + P.piHeI * c.HeI * 0.25 + P.piHeII * c.HeII * 0.25) * dom_inv;
    else if (P.iradshield == 1)
      edot = edot + ip * (P.piHI * c.HI * fH + P.piHeI * c.HeI * 0.25 + P.piHeII * c.HeII * 0.25) / dom;
    else if (P.iradshield == 2)
      edot = edot + ip * (P.piHI * c.HI * fH + P.piHeI * c.HeI * 0.25 * fHe + P.piHeII * c.HeII * 0.25) / dom;
    else if (P.iradshield == 3)
      edot = edot + ip * (P.piHI * c.HI * fH + P.piHeI * c.HeI * fHe) / dom;
  }

  if (PC == 0) {    // Cloudy primordial cooling + electron density from mu (:917-947)
    edot = edot + cloudy_edot(P, P.pri, th.rhoH, th.metallicity, logtem, 0, 0);
    double myde = 1 - th.mmw * (3.0 * P.fh + 1.0) / 4.0;
    if (P.imetal == 1) myde = myde - th.mmw * c.metal / (c.d * GRB_MU_METAL);
    myde = c.d * myde / th.mmw;
    th.myde = dmax(myde, 0.0);
  }

  if (P.igammah > 0) {   // photo-electric heating by dust (:951-1001)
    double geff;
    if (P.igammah == 1) {
      geff = (th.tgas > 2.e4) ? 0.0 : P.gammaha;
    } else if (P.igammah == 2) {
      geff = (th.tgas > 2.e4) ? 0.0 : P.gammaha * 0.05 * myisrf;
    } else {
      const double pe_X = myisrf * dom_inv * sqrt(th.tgas) / th.myde;
      const double pe_eps = (4.9e-2 / (1.0 + POW_POS(pe_X / 1925.0, 0.73))) +
                            ((3.7e-2 * POW_POS(th.tgas / 1.e4, 0.7)) / (1.0 + (pe_X / 5000.0)));
      geff = P.gammaha * pe_eps * myisrf;
    }
    edot = edot + geff * th.rhoH * dom_inv * th.dust2gas / P.fgr;
  }

  if ((P.idustall > 0) || (P.idustrec > 0)) {   // grain recombination cooling (:1005-1023)
    const double regr = cv(L_REGR);
    const double grbeta = 0.74 / POW_POS(th.tgas, 0.068);
    edot = edot - regr * POW_POS(myisrf * dom_inv / th.myde, grbeta) * th.myde * th.rhoH * th.dust2gas / P.fgr;
  }

  // Compton + X-ray Compton (:1027-1041)
  edot = edot - P.comp1 * (th.tgas - P.comp2) * th.myde * dom_inv -
         P.comp_xraya * (th.tgas - P.comp_temp) * th.myde * dom_inv;

  if (P.iradtrans == 1)
    edot = edot + (double)P.ipiht * c.photogamma / P.coolunit * c.HI / dom;

  if (P.imcool == 1) {   // Cloudy metal cooling (:1069-1097)
    if (P.clnew == 1) edot = edot + cloudy_edot(P, P.met, th.rhoH, th.metallicity, logtem, P.icmbTfloor, 1);
    else edot = edot + cloudy_old_edot(P, P.met, c.d, c.de, th.rhoH, th.metallicity, logtem);
  }

  if (P.iVheat == 1) edot = edot + c.vheat / P.coolunit / (dom * dom);
  if (P.iMheat == 1) edot = edot + c.mheat * c.d * GRB_MASS_H / P.coolunit / dom;

  tgasold = th.tgas;
  th.edot = edot;
}

// the whole of cool1d_multi_g for a kernel that evaluates it once per cell: the table
// values come straight from global memory, all loads of a column group in flight together
template <int PC>
__device__ inline void cool_rate(const GrParams &P, const Cell &c, bool first, double &tgasold,
                                 Thermo &th, bool lookup_follows = true)
{
  const double logtem = thermo_head<PC>(P, c, first, tgasold, th, lookup_follows);
  const bool anydust = (P.idust > 0) || (P.idustall > 0) || (P.idustrec > 0);
  CoolRegs cv;
#pragma unroll
  for (int i = 0; i < NCOOL; i++) cv.v[i] = 0.0;
  // the table index is only formed when species or the full dust model are on (:349-357);
  // otherwise the first bin is read as it stands
  const double *row = P.cool;
  double tdef = 0.0;
  if ((PC > 0) || (P.idustall > 0)) {
    const TIndex ti = t_index(P, logtem);
    row = P.cool + (size_t)(ti.idx - 1) * (NCOOL * 2);
    tdef = ti.tdef;
  }
  if (PC > 0) {
    double at[12];
    lerp_cols<L_CEHI, 12>(row, tdef, at);
#pragma unroll
    for (int i = 0; i < 12; i++) cv.v[L_CEHI + i] = at[i];
  }
  if (PC > 1) {
    double ga[6];
    lerp_cols<L_GAHI, 6>(row, tdef, ga);
#pragma unroll
    for (int i = 0; i < 6; i++) cv.v[L_GAHI + i] = ga[i];
    if (P.iciecool == 1) cv.v[L_CIECO] = lerp_row(row, L_CIECO, tdef);
  }
  if (PC > 2) {
    double hd[2];
    lerp_cols<L_HDLTE, 2>(row, tdef, hd);
    cv.v[L_HDLTE] = hd[0]; cv.v[L_HDLOW] = hd[1];
  }
  if (anydust) {
    cv.v[L_GASGR] = lerp_row(row, L_GASGR, tdef);
    cv.v[L_REGR] = lerp_row(row, L_REGR, tdef);
  }
  cool_sums<PC>(P, c, logtem, cv, tgasold, th);
}

}  // namespace grb
#endif
