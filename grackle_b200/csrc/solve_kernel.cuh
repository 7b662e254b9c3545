// grackle-b200: k_solve - solve_rate_cool_g.F (the whole file) for every active cell.
// Included by the per-network translation units solve_pc0.cu ... solve_pc3.cu (one per
// primordial_chemistry value, so the instantiations compile in parallel).
//
// Divergence control ("lane refill"): subcycle counts differ by orders of magnitude
// between cells, so a static thread<->cell mapping would leave most lanes of a warp idle
// while the slowest cell finishes (the reference's OpenMP rows have the same problem).
// Here a lane that finishes its cell immediately takes the next unprocessed cell from a
// global queue (one warp-aggregated atomicAdd per refill, found with __ballot_sync), so
// every lane executes useful subcycles until the queue is empty.  Each cell keeps its own
// iteration counter, hence the iter==1 / iter>50 heuristics of the reference are
// reproduced exactly and results do not depend on which lane processed a cell.
#ifndef GRB200_SOLVE_KERNEL_CUH
#define GRB200_SOLVE_KERNEL_CUH

#include "cell_io.cuh"

namespace grb {

// Per-cell carry between subcycles, and everything else a lane needs only now and then (the
// cell's field index, its metallicity and metal density - constant over the subcycles - and
// the lane's diagnostics).  All of it lives in the thread's own shared-memory columns behind the
// rate slots, not in registers: registers are what decides how many table loads the cooperative
// fetches keep in flight and whether the species update spills (measured, round 2: the carry
// alone, 9 registers, was worth 8 % of the kernel's time).
struct SubState {
  double &ttot, &tgasold, &dedot_prev, &HIdot_prev;
  int &iter;
};
enum { ST_TTOT = 0, ST_TGASOLD, ST_DEDOT, ST_HIDOT, ST_ITER, ST_X, ST_Z, ST_METAL, ST_DIAG0, ST_DIAG1, ST_NCOLS };

// One subcycle of solve_rate_cool_g.F:451-817 for one cell is split into per-cell stages
// with the warp-cooperative table fetches between them (see solve_cells): thermo_head, [fetch
// A], cool_sums + k13_martin, [fetch B], subcycle_lookup, subcycle_finish (rate-limited dt,
// energy update, species sweep).  subcycle_finish returns true when the cell has reached dt.
template <int PC, bool FOLD, class R>
__device__ __forceinline__ void subcycle_lookup(const GrParams &P, const Cell &c, const Thermo &th, size_t x,
                                                const TIndex &ti, const R &r)
{
  Neigh nb;
  if (PC > 1 && P.iH2shield == 1) {
    const double sc = (P.iexpand == 1) ? P.a3inv : 1.0;
    const size_t sj = (size_t)P.in, sk = (size_t)P.in * P.jn;
    const double *D = P.f[F_D];
    nb.d[0] = D[x + 1] * sc; nb.d[1] = D[x - 1] * sc; nb.d[2] = D[x + sj] * sc;
    nb.d[3] = D[x - sj] * sc; nb.d[4] = D[x + sk] * sc; nb.d[5] = D[x - sk] * sc;
  }
  lookup_rates<PC, FOLD>(P, c, th, nb, ti, r);
}

// k13dd: the density-dependent k13 of k13_martin (used when three_body_rate == 0)
template <int PC, bool FOLD, class R>
__device__ __forceinline__ bool subcycle_finish(const GrParams &P, Cell &c, SubState &s, const Thermo &th,
                                                const TIndex &ti, double k13dd, const R &r)
{
  const double dt = P.dt;
  double dtit = GRB_HUGE8;
  double edot = th.edot;
  double k13 = 0.0;
  if (PC > 1) k13 = (P.ithreebody == 0) ? k13dd : RK(k13);

  if (PC > 0) {
    double dedot, HIdot;
    rate_derivs<PC, FOLD>(P, c, th, r, k13, dedot, HIdot, edot);

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
    {
      FmFast f{P.fm_tab, false};
      double qe = f.div_s(0.1 * c.de, dedot), qh = f.div_s(0.1 * c.HI, HIdot);
      if (f.bad) { qe = ddiv_s(0.1 * c.de, dedot); qh = ddiv_s(0.1 * c.HI, HIdot); }
      dtit = dmin(dmin(fabs(qe), fabs(qh)), dmin(dt - s.ttot, 0.5 * dt));
    }

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
      const double heq = ddiv(-1.0, 4.0 * RK(k22)) * (k13 - sqrt(8.0 * k13 * RK(k22) * fd + k13 * k13));
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
  if (PC > 0) step_species<PC, FOLD>(P, c, th, r, k13, dtit, s.HIdot_prev, s.dedot_prev);

  s.ttot = dmin(s.ttot + dtit, dt);
  return fabs(dt - s.ttot) < 1.0e-10 * dt;
}

// extras that are only read when their physics is on are re-fetched per subcycle so they
// do not occupy registers across the loop
template <int PC>
__device__ __forceinline__ void load_extras(const GrParams &P, size_t x, Cell &c)
{
  load_cool_sources(P, x, c);
  load_solve_sources<PC>(P, x, c);
}

// CTA-wide re-alignment points.  The executed path of one subcycle (31 KB of SASS in the
// compact instantiation, 66 KB before round 1's code-size work) competes with a 32 KB
// instruction cache; warps that drift apart each stream it from L2 on their own (first
// version: 55 % of all stall cycles were stall_no_inst).  A barrier per subcycle keeps the
// warps of a group inside the same few KB of code, so one warp's fetch serves all of them.
// The CTA is split into lockstep groups of GW warps, each on its own named barrier.  With
// groups of 4 every SM sub-partition holds one warp of each group: the four warps a scheduler
// sees are in different phases and cover each other's latencies.
// The best group size depends on how much code a subcycle streams and on what else the warps wait
// for.  Round 1: 2 warps per group for the compact pc = 2/3 kernels, none for pc = 0/1, 4 for the
// general pc = 0/1 kernels, 8 for the general / dust + shielding kernels of the H2 networks.
// Round 2 (column-major slots, per-cell state in shared memory, fewer instructions): the compact
// kernels run fastest WITHOUT grouping (C4: 37.8 Mcell/s against 34.8 with groups of 2) - the
// barrier wait now costs more than the instruction-cache misses it avoids.  Re-measured at the end
// of round 2 (8 Mi cells, Set B, groups of 16 / 8 / 4 / 2 / 1 warps): general pc = 0 kernel - / - / 35.0 / 36.2 /
// 38.1 and general pc = 1 - / - / 74.2 / 80.8 / 88.7 Mcell/s (no grouping now); general pc = 2 19.3 / 19.7 / 16.3 /
// 13.8 / 14.1, general pc = 3 22.7 / 25.2 / 24.6 / 23.6 / 24.0, dust + shielding pc = 2 21.4 / 23.1 / 20.7 / 18.0 /
// 19.2 (groups of 8 stay).  The dust_chemistry kernel (VAR_DUST_ALL) prefers no grouping on uniform
// inputs (Set H: 11.6 against 10.6 Mcell/s) and groups of 8 on the broad Set B (23.2 against 22.4): 8 ships.
#ifndef GRB_GROUP_WARPS
#define GRB_GROUP_WARPS 1
#endif
#ifndef GRB_GROUP_WARPS_LEAN
#define GRB_GROUP_WARPS_LEAN 1
#endif
#ifndef GRB_GROUP_WARPS_LEAN_SMALL
#define GRB_GROUP_WARPS_LEAN_SMALL 1
#endif
#ifndef GRB_GROUP_WARPS_BIG
#define GRB_GROUP_WARPS_BIG 8
#endif
#ifndef GRB_GROUP_WARPS_DUSTALL
#define GRB_GROUP_WARPS_DUSTALL 8
#endif
static_assert(GRB_SOLVE_THREADS % (32 * GRB_GROUP_WARPS_BIG) == 0 && GRB_SOLVE_THREADS % (32 * GRB_GROUP_WARPS_DUSTALL) == 0,
              "CTA must be whole lockstep groups");
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

// ---- host-buffer pipeline hooks (GrParams::chunk_*; api.cu solve_pipelined) -----------
// They are a template switch of the kernel, not a run-time one: merely present in the
// subcycle loop they cost the device-resident path 2 % (measured), so that path runs
// instantiations without them.
// Per-lane state (the chunk of the cell in hand, the number of chunks last seen ready) lives
// in the lane's spare slot column, not in registers: the subcycle loop has none to give.
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

// COMPACT: the variant reads no dust / CIE cooling columns and produces no per-cell photo-rate
// slots (VAR_LEAN), so the slot columns shrink to slot_cols_needed (network.cuh)
template <int PC, bool FOLD, bool COMPACT, bool PIPE, int GW>
__device__ __forceinline__ void solve_cells(const GrParams &P)
{
  constexpr int G = coop_lanes(PC);
  constexpr int NCA = cool_cols(PC, COMPACT);             // cooling columns fetched in phase A
  constexpr int STATE_COL = slot_cols_needed(PC, COMPACT, PIPE) - 1;
  constexpr int CS = slot_col_stride(GRB_SOLVE_THREADS, G);
  extern __shared__ double slot_cols[];     // [slot_cols_needed][CS], column-major (network.cuh)
  double *const my_slots = slot_cols + threadIdx.x;
  double *const warp_slots = slot_cols + (threadIdx.x & ~31u);
  const Slots<CS> r{my_slots};
  const unsigned full = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const int bar_id = 1 + (int)(threadIdx.x >> 5) / GW;
  const bool anydust = (P.idust > 0) || (P.idustall > 0) || (P.idustrec > 0);
  Cell c;
  double *const st = my_slots + (size_t)slot_cols_needed(PC, COMPACT, PIPE) * CS;    // state columns behind the slots
  SubState s{st[ST_TTOT * CS], st[ST_TGASOLD * CS], st[ST_DEDOT * CS], st[ST_HIDOT * CS],
             *reinterpret_cast<int *>(st + ST_ITER * CS)};
  size_t &cell_x = *reinterpret_cast<size_t *>(st + ST_X * CS);
  // diagnostics: [0] subcycles, [1] most subcycles of one cell, [2] cells at max_iterations, [3] NaN cells
  unsigned *const my_diag = reinterpret_cast<unsigned *>(st + ST_DIAG0 * CS);     // [0], [1] here, [2], [3] one column on
  unsigned *const my_diag2 = reinterpret_cast<unsigned *>(st + ST_DIAG1 * CS);
  my_diag[0] = 0u; my_diag[1] = 0u; my_diag2[0] = 0u; my_diag2[1] = 0u;
  bool active = false, exhausted = false;
  // [0] chunk | unreported, [1] chunks seen ready
  unsigned *const lane_state = reinterpret_cast<unsigned *>(my_slots + (size_t)STATE_COL * CS);
  if (PIPE) { lane_state[0] = 0u; lane_state[1] = 0u; }

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
          // (the chunk-streaming path needs queue order = memory order: it runs without a permutation)
          const long long cell_t = (!PIPE && P.perm != nullptr) ? (long long)__ldg(P.perm + t) : t;
          const size_t x = cell_index(P, cell_t, ci, cj, ck);
          cell_x = x;
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
            load_cell<PC, false, true>(P, x, c, true);
            s.ttot = 0.0; s.tgasold = 0.0; s.dedot_prev = 0.0; s.HIdot_prev = 0.0; s.iter = 1;
            st[ST_METAL * CS] = c.metal;
            st[ST_Z * CS] = (P.imetal == 1) ? ddiv(c.metal, c.d) * P.inv_zsolar : 0.0;   // cool1d_multi_g.F:324-330
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
      load_extras<PC>(P, cell_x, c);
      c.metal = st[ST_METAL * CS];
      th.metallicity = st[ST_Z * CS];
      logtem = thermo_head<PC, false>(P, c, s.iter == 1, s.tgasold, th, true);
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
    if (PC > 1 && P.ithreebody == 0) {
      // the k13dd coefficients are read by k13_martin only (network.cuh: 500 K <= T < 1e6 K)
      const bool want_k13 = active && th.tgas >= 500.0 && th.tgas < 1.0e6;
      coop_fetch2<G, NCA, NCOOL * 16, 14, NCHEM * 16, CS>(P.cool, offA, tdefA, P.chem + 2 * C_K13DD0,
                                                          (GRB_SKIP_SEG2 && !want_k13) ? -1 : offB, tb.tdef,
                                                          warp_slots + SLA_COOL * CS);
    }
    else if (PC > 1 || anydust) coop_fetch<G, NCA, NCOOL * 16, CS>(P.cool, offA, tdefA, warp_slots + SLA_COOL * CS);
    else if (PC == 1) coop_fetch<G, L_BREM + 1, NCOOL * 16, CS>(P.cool, offA, tdefA, warp_slots + SLA_COOL * CS);
    __syncwarp();
    // stage 3 (per cell): cooling / heating sums, density-dependent k13
    double k13dd = GRB_TINY8;
    if (active) {
      cool_sums<PC>(P, c, logtem, CoolSlots<CS>{my_slots}, s.tgasold, th);
      if (PC > 1 && P.ithreebody == 0) k13dd = k13_martin<NCA>(P, c, th, r);
    }
    __syncwarp();
    if (PC > 0) {
      // stage 4 (warp): chemistry rows -> slots
      coop_fetch<G, chem_cols(PC), NCHEM * 16, CS>(P.chem, offB, tb.tdef, warp_slots);
      __syncwarp();
      // stage 5 (per cell): shielding, H2 on dust
      if (active) subcycle_lookup<PC, FOLD>(P, c, th, cell_x, tb, r);
    }
    if (active) {
      const bool done = subcycle_finish<PC, FOLD>(P, c, s, th, tb, k13dd, r);
      if (done || s.iter >= P.itmax) {
        const unsigned it = (unsigned)s.iter;        // = the subcycles this cell took
        my_diag[0] += it;
        if (it > my_diag[1]) my_diag[1] = it;
        if (!done) my_diag2[0] += 1u;
        if (!(c.e == c.e) || !(c.HI == c.HI)) my_diag2[1] += 1u;
        c.metal = st[ST_METAL * CS];
        store_cell<PC>(P, cell_x, c);
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
  unsigned long long w_sub = my_diag[0], w_max = my_diag[1], w_itmax = my_diag2[0], w_nan = my_diag2[1];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    w_sub += __shfl_xor_sync(full, w_sub, o);
    w_itmax += __shfl_xor_sync(full, w_itmax, o);
    w_nan += __shfl_xor_sync(full, w_nan, o);
    const unsigned long long m = __shfl_xor_sync(full, w_max, o);
    w_max = m > w_max ? m : w_max;
  }
  if (lane == 0) {
    atomicAdd(P.diag + 0, w_sub);
    atomicMax(P.diag + 1, w_max);
    if (w_itmax) atomicAdd(P.diag + 2, w_itmax);
    if (w_nan) atomicAdd(P.diag + 3, w_nan);
  }
}

// dynamic shared memory of one CTA: the slot columns + the log/exp table
template <int PC, int VAR, bool PIPE>
__host__ __device__ constexpr size_t solve_smem_doubles()
{
  return (size_t)(slot_cols_needed(PC, VAR == VAR_LEAN, PIPE) + ST_NCOLS) * slot_col_stride(GRB_SOLVE_THREADS, coop_lanes(PC));
}
template <int PC, int VAR, bool PIPE>
__host__ __device__ constexpr size_t solve_smem_bytes()
{
  return (solve_smem_doubles<PC, VAR, PIPE>() + GRB_FM_TABLE_DOUBLES) * sizeof(double);
}

// lockstep group size of an instantiation (see above)
template <int PC, int VAR>
__host__ __device__ constexpr int solve_group_warps()
{
  return (VAR == VAR_LEAN) ? ((PC <= 1) ? GRB_GROUP_WARPS_LEAN_SMALL : GRB_GROUP_WARPS_LEAN)
                           : ((PC <= 1) ? GRB_GROUP_WARPS : (VAR == VAR_DUST_ALL ? GRB_GROUP_WARPS_DUSTALL : GRB_GROUP_WARPS_BIG));
}

template <int PC, int VAR, bool PIPE>
__global__ void __launch_bounds__(GRB_SOLVE_THREADS, GRB_SOLVE_MINBLOCKS) k_solve(const __grid_constant__ GrParams P0)
{
  // the 2.5 KB log/exp table goes to shared memory behind the slot columns; the by-value copy of
  // the parameters is what lets the compiler fold the variant's switches and the table pointer
  extern __shared__ double slot_cols[];
  double *const fm = slot_cols + solve_smem_doubles<PC, VAR, PIPE>();
  for (int i = threadIdx.x; i < GRB_FM_TABLE_DOUBLES; i += GRB_SOLVE_THREADS) fm[i] = P0.fm_tab[i];
  __syncthreads();
  GrParams P = P0;
  P.fm_tab = fm;
  variant_flags<VAR>(P);
  solve_cells<PC, variant_folds(VAR), VAR == VAR_LEAN, PIPE, solve_group_warps<PC, VAR>()>(P);
}

// ---- per-network launchers (one translation unit each) ---------------------------------
template <int PC, int VAR, bool PIPE>
inline void launch_solve_inst(const GrParams &P, int blocks, cudaStream_t st)
{
  constexpr size_t smem = solve_smem_bytes<PC, VAR, PIPE>();
  cudaFuncSetAttribute(k_solve<PC, VAR, PIPE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  k_solve<PC, VAR, PIPE><<<blocks, GRB_SOLVE_THREADS, smem, st>>>(P);
}

template <int PC>
inline void launch_solve_pc(const GrParams &P, int var, int blocks, cudaStream_t st)
{
#ifdef GRB_DEV_ONLY_LEAN      /* development builds: only the compact device-resident instantiation */
  launch_solve_inst<PC, VAR_LEAN, false>(P, blocks, st);
#else
#define GRB_LAUNCH_VAR(V)                                        \
  {                                                              \
    if (P.chunk_shift > 0) launch_solve_inst<PC, V, true>(P, blocks, st);   \
    else launch_solve_inst<PC, V, false>(P, blocks, st);         \
  }
  if (var == VAR_LEAN) GRB_LAUNCH_VAR(VAR_LEAN)
  else if (var == VAR_DUST_SHIELD) GRB_LAUNCH_VAR(VAR_DUST_SHIELD)
  else if (var == VAR_DUST_ALL) GRB_LAUNCH_VAR(VAR_DUST_ALL)
  else if (var == VAR_COMMON) GRB_LAUNCH_VAR(VAR_COMMON)
  else if (var == VAR_RANK45) GRB_LAUNCH_VAR(VAR_RANK45)
  else GRB_LAUNCH_VAR(VAR_GENERAL)
#undef GRB_LAUNCH_VAR
#endif
}

template <int PC>
inline int solve_max_blocks_pc()
{
  int n = 0;
#ifdef GRB_DEV_ONLY_LEAN
  auto kern = k_solve<PC, VAR_LEAN, false>;
  constexpr size_t smem = solve_smem_bytes<PC, VAR_LEAN, false>();
#else
  auto kern = k_solve<PC, VAR_GENERAL, true>;
  constexpr size_t smem = solve_smem_bytes<PC, VAR_GENERAL, true>();
#endif
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kern, GRB_SOLVE_THREADS, smem);
  return n;
}

}  // namespace grb
#endif
