// optics_rorr_cut.cu -- k-interpolation + random overlap with resorting and rebinning WITHOUT a sort.
//
// Replaces, for a batch of columns, the body of the per-bin loop of
// OpticalProperties_compute_opacity (src/radtran/clima_radtran_types.f90:638-771) and k_rorr
// (:780-888), like rorr_impl.cuh, but computes the rebinned mixture from "staircase cuts":
//
//   k_rorr ranks the G*G sums tau_xy = a_i + b_j (mrgrnk, :840), accumulates the weights
//   w_i*w_j in sorted order (weights_to_bins, :843) and rebins onto the G original g-bins (futils
//   rebin, :847).  The result only depends on the running integral of the sorted sequence,
//   I(E) = int_0^E tau(c) dc, at the G-1 interior bin edges:  new(g) = (I(E_g+1) - I(E_g)) / dE_g.
//   For one edge E,   I(E) = S(cut) + tau_e * (E - W(cut)),   where `cut` is the set of elements
//   sorted before the element e that contains E.  Because a and b ascend, a cut is a staircase
//   c_i = number of elements of row i inside it, and with Wb[c] = sum_{j<c} w_j,
//   Sb[c] = sum_{j<c} w_j b_j:
//        W(cut) = sum_i w_i Wb[c_i]          S(cut) = sum_i w_i (a_i Wb[c_i] + Sb[c_i]).
//   A staircase IS the sorted prefix that contains E iff
//        L = max_i a_i + b[c_i - 1]  <=  R = min_i a_i + b[c_i]      (a down-set of the order)
//        W <= E < W + w_e                                            (e = the argmin of R)
//   -- an exact test on the exact sums: no key quantisation, no certificate, no second kernel.
//   Ties need no order: I(E) depends on the distribution of the sums only.
//
// One lane per bin edge (lane G-1: the whole integral), P lanes per unit, 32/P units per warp.  One
// PASS over the G rows evaluates L, R, W, S for the lane's staircase; a staircase that fails the
// test is repaired by single-element moves (drop the largest inside element / take the smallest
// outside one), after 8 of them by jumps over whole groups of tied sums, and re-evaluated.  The
// staircases of a unit are the initial guess of the NEXT unit of a CHAIN: the same (bin, layer slot)
// of the next column when neighbouring columns are alike (a sweep: 1.06-1.19 passes per merge; the
// default use), or the next layer slot of the same (column, bin) (2.3-2.5 passes per merge: slower than
// the sorting kernels, kept behind CLIMA_RT_RORR_CHAIN=slot; experiments/rorr_cut_proto.py).
// The first unit of a chain starts from a bisection on the IEEE-754 pattern of the edge element's
// value (until <= 3 elements lie between the brackets).  Merges whose rows or columns do not
// interleave take the closed form a_i + <b> / b_j + <a>.  A unit that does not settle within the
// pass budget (none in any measured workload) is appended to the worklist of the exact sorting
// kernel (optics_rorr_exact.cu).
//
// What it costs (ncu, profiles/r2b_ncu_rorr_cut_summary.txt): 4 shared-memory gathers + ~24
// instructions per (edge, row) and pass, 15 x 16 pairs per merge; 16.7 M warp-instructions per column
// against 24.2 M for the sorting kernel; shared-memory wavefronts 69 % of peak, issue slots 66 %.
// DESIGN.md section 3.1 has the derivation, the measurements and the dead ends.
#include <cstdio>
#include <cstdlib>

#include "rt_device.cuh"
#include "rt_kernels.h"

#ifndef CUT_WPB
#define CUT_WPB 4  // warps per block
#endif
#ifndef CUT_MIN_BLOCKS
#define CUT_MIN_BLOCKS 6  // 80 registers, 24 warps per SM
#endif
#ifndef CUT_UNIT_PAD
#define CUT_UNIT_PAD 0
#endif
#ifndef CUT_ROW_UNROLL
#define CUT_ROW_UNROLL 2  // count words (4 rows) per trip of the pass loop: fully unrolled it needs 168 registers.  r2g, 4096 cfg2
                          // columns: 1 -> 85.1 ms, 2 -> 83.2 ms (150 bytes of spills at 80 registers), 4 -> 95-140 ms at 3-5 blocks per SM
#endif

namespace rorr_cut {

template <int P>
struct Cfg {
  static constexpr int LOGP = (P == 8 ? 3 : (P == 16 ? 4 : 5));
  static constexpr int UPW = 32 / P;        // units per warp
  static constexpr int UPB = CUT_WPB * UPW;  // chains per block
  static constexpr int CWN = P / 4;          // count words (4 rows each) per lane and merge
  static constexpr int ROWU = CUT_ROW_UNROLL;
  // The pass gathers four numbers by the row's count c in 0..P.  The lanes of a unit (one per edge)
  // hold non-decreasing counts for a given row, typically a run of 0 (row outside), a few values in
  // between, and a run of P (row inside): with 8-byte entries c = 0 and c = P fall into the same
  // banks and nearly every gather took two wavefronts (ncu r2a: x1.8).  Counts are therefore kept as
  // POSITIONS pos(c) = c for c < P, P + P/2 for c = P: the full row shares its banks with c = P/2.
  static constexpr int NPOS = P + P / 2 + 1;
  // per-unit tables, in doubles
  static constexpr int BX = 0;               // search table: [0] = -inf, [1..G] = b, [G+1..2P+1] = +inf
  static constexpr int BM = 2 * P + 2;       // by position: b[c-1]  (-inf for c = 0)
  static constexpr int BP = BM + NPOS;       //              b[c]    (+inf for c = G)
  static constexpr int SB = BP + NPOS;       //              Sb[c] = sum_{j<c} w_j b_j
  static constexpr int WB = SB + NPOS;       //              Wb[c] = sum_{j<c} w_j   (static)
  static constexpr int AT = WB + NPOS;       // a[0..P-1]
  static constexpr int TT = AT + P;          // scratch: total tau of the unit (band sum)
  static constexpr int UNIT = TT + P + CUT_UNIT_PAD;  // doubles per unit (the pad sets the bank offset between the units of a warp)
  static constexpr int NBIS = 44;            // cap on the bisection steps of a cold start (a plateau of nearly equal rows needs ~25 to get inside it)
  static constexpr int MAXPASS = 48;         // passes before a restart from bisection / giving up
  static constexpr int SINGLES = 8;          // single-element moves of a lane before it jumps by tie groups
  __host__ __device__ static constexpr unsigned pos_of(unsigned c) { return c + ((c >> LOGP) << (LOGP - 1)); }
  __host__ __device__ static constexpr unsigned cnt_of(unsigned pb) { return pb < (unsigned)P ? pb : (unsigned)P; }
};

__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ double lds_f64(unsigned addr) {
  double v;
  asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ void sts_f64(unsigned addr, double v) {
  asm volatile("st.shared.f64 [%0], %1;" ::"r"(addr), "d"(v) : "memory");
}
__device__ __forceinline__ unsigned lds_u32(unsigned addr) {
  unsigned v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ void sts_u32(unsigned addr, unsigned v) {
  asm volatile("st.shared.u32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}

struct PassOut {
  double L, R, W, S;
  int iL, iR;  // rows of the largest element inside / the smallest element outside
};

// One pass over the rows for the lane's staircase (position bytes at cnt, 4 rows per word, word
// stride 128 bytes).  Four rows per trip; inside a trip the extrema and the sums are formed as
// trees / from zero, so that only one compare-select and one add per trip sit on the loop-carried
// dependency chains (the pass is latency bound: r2a 56 % issue slots, stall_wait 2.7 per issue).
template <int P, bool FULL>
__device__ __forceinline__ PassOut cut_pass(const unsigned ua, const unsigned cnt, const double (&w)[32], const int G) {
  using C = Cfg<P>;
  const double NINF = __longlong_as_double(0xfff0000000000000ll), PINF = __longlong_as_double(0x7ff0000000000000ll);
  PassOut o;
  o.L = NINF;
  o.R = PINF;
  o.W = 0.0;
  o.S = 0.0;
  o.iL = 0;
  o.iR = 0;
#pragma unroll(C::ROWU)
  for (int q = 0; q < C::CWN; q++) {
    const unsigned cw = lds_u32(cnt + q * 128u);
    double lo[4], hi[4], t1[4], wb[4];
#pragma unroll
    for (int r = 0; r < 4; r++) {
      const int i = 4 * q + r;
      const unsigned pb = (cw >> (8 * r)) & 0xffu;
      const unsigned ad = ua + pb * 8u;
      const double bm = lds_f64(ad + C::BM * 8u), bp = lds_f64(ad + C::BP * 8u);
      const double sb = lds_f64(ad + C::SB * 8u);
      wb[r] = lds_f64(ad + C::WB * 8u);
      const double ai = lds_f64(ua + (unsigned)(C::AT + i) * 8u);
      lo[r] = __dadd_rn(ai, bm);  // :828 tau_xy of the last element inside ...
      hi[r] = __dadd_rn(ai, bp);  // ... and of the first one outside
      t1[r] = __fma_rn(ai, wb[r], sb);
      if (!FULL && i >= G) { lo[r] = NINF; hi[r] = PINF; t1[r] = 0.0; wb[r] = 0.0; }
    }
    {  // largest inside / smallest outside of the trip (ties: any), then against the running ones
      const bool p01 = lo[1] > lo[0], p23 = lo[3] > lo[2];
      const double l01 = p01 ? lo[1] : lo[0], l23 = p23 ? lo[3] : lo[2];
      const int i01 = p01 ? 1 : 0, i23 = p23 ? 3 : 2;
      const bool pq = l23 > l01;
      const double lq = pq ? l23 : l01;
      const int iq = pq ? i23 : i01;
      if (lq > o.L) { o.L = lq; o.iL = 4 * q + iq; }
    }
    {
      const bool p01 = hi[1] < hi[0], p23 = hi[3] < hi[2];
      const double h01 = p01 ? hi[1] : hi[0], h23 = p23 ? hi[3] : hi[2];
      const int i01 = p01 ? 1 : 0, i23 = p23 ? 3 : 2;
      const bool pq = h23 < h01;
      const double hq = pq ? h23 : h01;
      const int iq = pq ? i23 : i01;
      if (hq < o.R) { o.R = hq; o.iR = 4 * q + iq; }
    }
    // w[i]: constant-bank operands (DevTables::wq)
    double sq = __dmul_rn(w[4 * q], t1[0]), wq = __dmul_rn(w[4 * q], wb[0]);
#pragma unroll
    for (int r = 1; r < 4; r++) {
      sq = __fma_rn(w[4 * q + r], t1[r], sq);
      wq = __fma_rn(w[4 * q + r], wb[r], wq);
    }
    o.S = __dadd_rn(o.S, sq);
    o.W = __dadd_rn(o.W, wq);
  }
  return o;
}

// position byte of row i of the staircase at cnt
__device__ __forceinline__ unsigned row_pos(const unsigned cnt, const int i) {
  return (lds_u32(cnt + (unsigned)(i >> 2) * 128u) >> (8 * (i & 3))) & 0xffu;
}
__device__ __forceinline__ void row_set(const unsigned cnt, const int i, const unsigned pb) {
  const unsigned a = cnt + (unsigned)(i >> 2) * 128u;
  const int sh = 8 * (i & 3);
  sts_u32(a, (lds_u32(a) & ~(0xffu << sh)) | (pb << sh));
}

// Staircase of all elements with a_i + b_j < X (LE: <= X), compared on the rounded sums the pass
// uses; returns its weight W and the packed positions.  Out of line: repair path only.
template <int P>
struct Stair {
  unsigned cw[Cfg<P>::CWN];
  double W;
  int n;  // elements inside
};
template <int P, bool LE>
__device__ __noinline__ Stair<P> stair_of(const unsigned ua, const unsigned swa, const int G, const double X) {
  using C = Cfg<P>;
  Stair<P> st;
  st.W = 0.0;
  st.n = 0;
#pragma unroll 1
  for (int q = 0; q < C::CWN; q++) {
    unsigned cw = 0;
#pragma unroll
    for (int r = 0; r < 4; r++) {
      const int i = 4 * q + r;
      if (i < G) {
        const double ai = lds_f64(ua + (unsigned)(C::AT + i) * 8u);
        unsigned c = 0;
#pragma unroll
        for (int s = P; s >= 1; s >>= 1) {
          const double v = __dadd_rn(ai, lds_f64(ua + (C::BX + c + s) * 8u));
          if (LE ? (v <= X) : (v < X)) c += s;
        }
        const unsigned pb = C::pos_of(c);
        cw |= pb << (8 * r);
        st.n += (int)c;
        st.W = __fma_rn(lds_f64(swa + (unsigned)i * 8u), lds_f64(ua + (C::WB + pb) * 8u), st.W);
      }
    }
    st.cw[q] = cw;
  }
  return st;
}
template <int P>
__device__ __forceinline__ void stair_store(const unsigned cnt, const Stair<P> &st) {
#pragma unroll
  for (int q = 0; q < Cfg<P>::CWN; q++) sts_u32(cnt + q * 128u, st.cw[q]);
}

// Cold start: bisection (on the IEEE-754 pattern) for the value of the element that contains edge
// E, until at most 3 elements lie between the brackets (or NBIS steps), then the staircase of everything below
// the lower bracket.  Only an initial guess: cut_pass decides.  Out of line and rolled: it runs once
// per chain (and merge), the pass loop must stay small.
template <int P>
__device__ __noinline__ void cut_bisect(const unsigned ua, const unsigned cnt, const unsigned swa,
                                        const int G, const double E, const double tol) {
  using C = Cfg<P>;
  long long lo = __double_as_longlong(__dadd_rn(lds_f64(ua + C::AT * 8u), lds_f64(ua + (C::BX + 1) * 8u)));
  long long hi = __double_as_longlong(__dadd_rn(lds_f64(ua + (unsigned)(C::AT + G - 1) * 8u), lds_f64(ua + (unsigned)(C::BX + G) * 8u)));
  int nlo = 0, nhi = G * G;
  // plain bisection: regula falsi on (pattern, cumulative weight) was tried and is slower (26 against
  // 17 steps on average: the weight below t is a staircase with long flats in the pattern)
#pragma unroll 1
  for (int it = 0; it < C::NBIS && hi - lo > 1 && nhi - nlo > 3; it++) {
    const double t = __longlong_as_double(lo + ((hi - lo) >> 1));
    const Stair<P> st = stair_of<P, false>(ua, swa, G, t);
    if (st.W > E + tol) { hi = __double_as_longlong(t); nhi = st.n; } else { lo = __double_as_longlong(t); nlo = st.n; }
  }
  stair_store<P>(cnt, stair_of<P, false>(ua, swa, G, __longlong_as_double(lo)));
}

// k(g) * column amount of k-gas jj for one unit: bilinear log10k(P,T) interpolation, lane per
// g-point (A.2, :649-662 + interp_2d, linear_interpolation_module.F90:322-335).  Split in two so
// that the table loads are in flight while the previous merge runs.
struct KRaw {
  double f0, f1, f2, f3, q1, q2, col;
};
struct UnitIn {  // per-unit bases of the k-gas inputs (64-bit address arithmetic once per unit, not per gas)
  const int *ib;
  const double *qb, *dens;
  double dz;
};
__device__ __forceinline__ UnitIn unit_in(const DevTables &tb, const DevChunk &ch, const int c, const int slot, const int k) {
  const int nz = ch.nz;
  const size_t rec = ((size_t)c * nz + slot) * rt_nbr(tb);
  UnitIn u;
  u.ib = ch.ibr + rec;
  u.qb = ch.qbr + rec;
  u.dens = ch.dens + (size_t)c * nz * tb.ng + k;
  u.dz = ch.dz[(size_t)c * nz + k];
  return u;
}
__device__ __forceinline__ KRaw kgas_load(const DevTables &tb, const UnitIn &u, const int nz, const int l,
                                          const int jj, const int g) {
  const int G = tb.G;
  const DevK &kt = tb.k[jj];
  const int ix = u.ib[2 * jj], iy = u.ib[2 * jj + 1];
  KRaw r;
  r.q1 = u.qb[2 * jj];
  r.q2 = u.qb[2 * jj + 1];
  r.col = u.dens[(unsigned)((kt.sp - 1) * nz)] * u.dz;  // cols(j,sp) :607
  const int gc = g < G ? g : G - 1;
  const int sy = G * kt.np;
  const double *f = kt.log10k + (unsigned)(G * (ix + kt.np * (iy + kt.nt * l)) + gc);
  r.f0 = f[0];
  r.f1 = f[G];
  r.f2 = f[sy];
  r.f3 = f[sy + G];
  return r;
}
__device__ __forceinline__ double kgas_finish(const KRaw &r, const bool live) {
  const double p1 = 1.0 - r.q1, p2 = 1.0 - r.q2;
  const double fx1 = p1 * r.f0 + r.q1 * r.f1;
  const double fx2 = p1 * r.f2 + r.q1 * r.f3;
  const double lk = p2 * fx1 + r.q2 * fx2;
  const double kv = (lk > -280.0 && lk < 280.0) ? rt_ten2power_lean(lk) : rt_ten2power_cold(lk);
  return live ? kv * r.col : 0.0;  // :660, :816-819 / :828
}

// mode 0: a chain = all unique layer slots of one (column, bin); the units of a warp are adjacent bins.
// mode 1: a chain = the same (bin, slot) of `Lc` consecutive columns; the units of a warp are adjacent slots.
template <int P, bool FULL>
__global__ void __launch_bounds__(32 * CUT_WPB, CUT_MIN_BLOCKS)
rorr_cut_kernel(const __grid_constant__ DevTables tb, const __grid_constant__ DevChunk ch,
                const int mode, const int Lc,
                int *__restrict__ worklist, int *__restrict__ wcount, const int wcap) {
  using C = Cfg<P>;
  constexpr double TOL = 0x1p-48;  // an edge that coincides with an element boundary: W and E differ by rounding only
  extern __shared__ double s_dyn[];
  __shared__ double s_w[32];
  const int G = FULL ? P : tb.G, nw = tb.nw, nz = ch.nz, cap = ch.nslot_cap, nk = tb.nk;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int ul = lane & (P - 1);
  const int ub = threadIdx.x / P;
  double *ut = s_dyn + (size_t)ub * C::UNIT;
  const unsigned ua = smem_u32(ut), swa = smem_u32(s_w);
  const unsigned cnt0 = smem_u32(reinterpret_cast<unsigned *>(s_dyn + (size_t)C::UPB * C::UNIT) +
                                 (size_t)warp * (nk - 1) * C::CWN * 32 + lane);
  const unsigned umask = (P == 32 ? 0xffffffffu : ((1u << P) - 1u)) << (lane & ~(P - 1));
  const bool isEdge = ul < G - 1, isTot = ul == G - 1, live = ul < G;
  const double E = isEdge ? tb.wbin_e[ul + 1] : 0.0;
  const double wl = live ? tb.wbin[ul] : 0.0, idw = live ? tb.inv_dw[ul] : 0.0;

  // ---- static parts of the unit tables and of the staircases
  if (threadIdx.x < 32) s_w[threadIdx.x] = threadIdx.x < G ? tb.wbin[threadIdx.x] : 0.0;
  {
    const double NINF = __longlong_as_double(0xfff0000000000000ll), PINF = __longlong_as_double(0x7ff0000000000000ll);
    for (int i = ul; i < C::UNIT; i += P) {
      double v = 0.0;
      if (i < C::BM) v = (i == 0) ? NINF : PINF;                  // search table sentinels
      else if (i < C::BP) v = NINF;                               // b[c-1]: position 0 stays -inf
      else if (i < C::SB) v = PINF;                               // b[c]: position of c = G stays +inf
      ut[i] = v;
    }
    __syncwarp();
    for (int c = ul; c <= G; c += P) ut[C::WB + C::pos_of((unsigned)c)] = tb.wbin_e[c];
    const unsigned full = C::pos_of((unsigned)G) * 0x01010101u;
    for (int m = 0; m < nk - 1; m++)
      for (int q = 0; q < C::CWN; q++) {
        unsigned wq = 0;
        if (isTot) {  // the whole integral: every row complete
          wq = full;
          if (!FULL)
            for (int r = 0; r < 4; r++)
              if (4 * q + r >= G) wq &= ~(0xffu << (8 * r));
        }
        sts_u32(cnt0 + (unsigned)(m * C::CWN + q) * 128u, wq);
      }
  }
  __syncthreads();

  // ---- which chain
  const unsigned id = blockIdx.x * (unsigned)C::UPB + (unsigned)ub;
  int c, l, slot, len;
  bool chain_ok;
  if (mode == 0) {
    const unsigned nwp = (unsigned)((nw + C::UPW - 1) / C::UPW * C::UPW);
    const unsigned cc = id / nwp;
    l = (int)(id - cc * nwp);
    chain_ok = cc < (unsigned)ch.ncol && l < nw;
    c = cc < (unsigned)ch.ncol ? (int)cc : ch.ncol - 1;
    if (l >= nw) l = nw - 1;
    slot = 0;
    len = ch.nslots[c];
  } else {
    const unsigned capp = (unsigned)((cap + C::UPW - 1) / C::UPW * C::UPW);
    const unsigned q1 = id / capp;
    slot = (int)(id - q1 * capp);
    const unsigned seg = q1 / (unsigned)nw;
    l = (int)(q1 - seg * (unsigned)nw);
    const long long c0 = (long long)seg * Lc;
    chain_ok = c0 < ch.ncol && slot < cap;
    c = c0 < ch.ncol ? (int)c0 : ch.ncol - 1;
    len = chain_ok ? (ch.ncol - c < Lc ? ch.ncol - c : Lc) : 0;
    len = __shfl_sync(0xffffffffu, len, 0);  // the units of a warp share (seg, l); unit 0 always has slot < cap if any has
    if (slot >= cap) slot = cap - 1;
  }
  if (__all_sync(0xffffffffu, !chain_ok)) return;

  unsigned need_cold = 0xffffffffu;  // merges whose staircases have not been initialised yet (bit jj)
  unsigned st_merges = 0, st_passes = 0, st_restarts = 0, st_fail = 0, st_cold = 0, st_closed = 0;
  for (int n = 0; n < len; n++) {
    bool active = chain_ok;
    int us = slot;  // the unit's slot
    if (mode == 0) {
      us = n;
    } else {
      if (n > 0) c++;
      const int nsl = ch.nslots[c];
      if (us >= nsl) { us = nsl - 1; active = false; }
    }
    const int k = ch.layer_of_slot[(size_t)c * nz + us];  // ground-first layer index
    const size_t obase = (((size_t)c * nw + l) * cap + us);
    const size_t U = (size_t)ch.ncol * nw * cap;
    const double q_s3 = ch.cont[obase], q_tauc = ch.cont[U + obase], q_sc = ch.cont[2 * U + obase];
    bool unit_ok = true;

    const UnitIn uin = unit_in(tb, ch, c, us, k);
    KRaw raw = kgas_load(tb, uin, nz, l, 0, ul);
    for (int jj = 0; jj < nk; jj++) {
      const double v = kgas_finish(raw, live);
      if (jj + 1 < nk) raw = kgas_load(tb, uin, nz, l, jj + 1, ul);
      if (jj == 0) {
        if (live) sts_f64(ua + (unsigned)(C::AT + ul) * 8u, v);  // tau_k(:,g) = k_1 * col_1  :816-819
        __syncwarp();
        continue;
      }
      // ---- tables of the new gas: b (with sentinels) and Sb[c] = sum_{j<c} w_j b_j
      double s = __dmul_rn(wl, v);
#pragma unroll
      for (int d = 1; d < P; d <<= 1) {
        const double t = __shfl_up_sync(0xffffffffu, s, d, P);
        if (ul >= d) s = __dadd_rn(s, t);
      }
      if (live) {
        const unsigned p1 = C::pos_of((unsigned)ul + 1u);
        sts_f64(ua + (unsigned)(C::BX + 1 + ul) * 8u, v);
        sts_f64(ua + (unsigned)(C::BP + ul) * 8u, v);  // pos_of(ul) = ul
        sts_f64(ua + (C::BM + p1) * 8u, v);
        sts_f64(ua + (C::SB + p1) * 8u, s);
      }
      __syncwarp();
      // ---- merges that need no staircase.  If the rows do not interleave (a_i + b_G <= a_i+1 + b_1
      // for every i) the sorted order is the row-major index order, row i covers exactly the
      // cumulative-weight interval of new bin i, and the rebin (:843-849) returns
      // a_i + sum_j w_j b_j = a_i + <b>; symmetrically b_j + <a> when the columns do not interleave.
      // The decision is per unit (a qualifying unit takes the closed form also when its warp evaluates
      // staircases for its other unit); the unit's staircases are left as they are for its successor.
      bool triv;
      double cf = 0.0;
      {
        const unsigned gn = (ul + 1 < G) ? ul + 1 : ul;
        const double a1 = lds_f64(ua + C::AT * 8u), aG = lds_f64(ua + (unsigned)(C::AT + G - 1) * 8u);
        const double am = lds_f64(ua + (unsigned)(C::AT + ul) * 8u), an = lds_f64(ua + (C::AT + gn) * 8u);
        const double b1 = lds_f64(ua + (C::BX + 1) * 8u), bG = lds_f64(ua + (unsigned)(C::BX + G) * 8u);
        const double bn = lds_f64(ua + (C::BX + 1 + gn) * 8u);
        const bool rowsep = !(ul < G - 1) || (__dadd_rn(am, bG) <= __dadd_rn(an, b1));
        const bool colsep = !(ul < G - 1) || (__dadd_rn(v, aG) <= __dadd_rn(bn, a1));
        const unsigned br = __ballot_sync(0xffffffffu, rowsep), bc = __ballot_sync(0xffffffffu, colsep);
        const bool urow = (br & umask) == umask, ucol = (bc & umask) == umask;
        triv = !(tb.pad0 & 1) && (urow || ucol);
        if (__any_sync(0xffffffffu, triv)) {
          const double bmean = lds_f64(ua + (C::SB + C::pos_of((unsigned)G)) * 8u);  // <b> = Sb[G]
          double amean = 0.0;
          if (__any_sync(0xffffffffu, triv && !urow)) {  // <a>, summed like Sb; rarely needed
            amean = __dmul_rn(wl, am);
#pragma unroll
            for (int d = 1; d < P; d <<= 1) {
              const double t = __shfl_up_sync(0xffffffffu, amean, d, P);
              if (ul >= d) amean = __dadd_rn(amean, t);
            }
            amean = __shfl_sync(0xffffffffu, amean, G - 1, P);
          }
          cf = urow ? __dadd_rn(am, bmean) : __dadd_rn(v, amean);
        }
      }
      if (__all_sync(0xffffffffu, triv)) {
        __syncwarp();
        if (live) sts_f64(ua + (unsigned)(C::AT + ul) * 8u, cf);
        __syncwarp();
        st_closed++;
        continue;
      }
      const unsigned cntm = cnt0 + (unsigned)((jj - 1) * C::CWN) * 128u;
      const bool cold = (need_cold >> jj) & 1u;
      if (cold) {
        if (isEdge) cut_bisect<P>(ua, cntm, swa, G, E, TOL);
        need_cold &= ~(1u << jj);
        st_cold++;
      }
      st_merges++;
      // ---- evaluate / repair the staircases
      PassOut o;
      bool ok, tie_ok = false;
      int passes = 0, moves = 0;
      bool restarted = false;
      while (true) {
        o = cut_pass<P, FULL>(ua, cntm, tb.wq, G);
        st_passes++;
        const unsigned pbR = row_pos(cntm, o.iR);
        const double we = __dmul_rn(s_w[o.iR], s_w[pbR < (unsigned)G ? pbR : G - 1]);  // weight of the element that would come next
        const bool inv = o.L > o.R;
        const bool over = o.W > E + TOL;
        const bool under = E >= __dadd_rn(o.W, we) - TOL;  // an edge AT an element boundary belongs to the staircase that ends there
        ok = !isEdge || tie_ok || triv || !(inv || over || under);
        if (__all_sync(0xffffffffu, ok)) break;
        if (++passes >= C::MAXPASS) {
          if (restarted) break;
          restarted = true;
          st_restarts++;
          passes = 0;
          if (!ok) cut_bisect<P>(ua, cntm, swa, G, E, TOL);
          continue;
        }
        if (!ok) {
          // single-element moves first (the staircase of the chain predecessor is a few elements
          // off); a lane that keeps moving is inside a group of tied sums and jumps by groups
          const bool single = moves++ < C::SINGLES;
          if (inv || (single && over)) {  // drop the largest element inside
            const unsigned pbL = row_pos(cntm, o.iL);
            row_set(cntm, o.iL, C::pos_of(C::cnt_of(pbL) - 1u));
          }
          if (inv || (single && !over && under)) {  // take the smallest element outside
            row_set(cntm, o.iR, C::pos_of(C::cnt_of(row_pos(cntm, o.iR)) + 1u));
          }
          if (!inv && !single) {
            if (over) {  // drop the largest element inside and everything tied with it
              stair_store<P>(cntm, stair_of<P, false>(ua, swa, G, o.L));
            } else {  // take the next element and everything tied with it ...
              const Stair<P> st = stair_of<P, true>(ua, swa, G, o.R);
              // ... unless the edge falls INSIDE that tie group: every element between the staircase
              // and the edge then has the value R, and I(E) = S + R (E - W) holds as it stands
              if (E < st.W - TOL) tie_ok = true; else stair_store<P>(cntm, st);
            }
          }
        }
      }
      {
        const unsigned bal = __ballot_sync(0xffffffffu, ok);
        unit_ok = unit_ok && ((bal & umask) == umask);
      }
#ifdef CUT_DEBUG
      if (!ok && atomicAdd(ch.cut_stats + 7, 1ull) < 4ull) {
        printf("FAIL c=%d l=%d slot=%d jj=%d edge=%d E=%.17g L=%.17g iL=%d R=%.17g iR=%d W=%.17g S=%.17g\n", c, l, us, jj, ul, E, o.L, o.iL, o.R, o.iR, o.W, o.S);
        for (int i = 0; i < G; i++)
          printf("  e%d row %2d a=%.17g b=%.17g pos=%u\n", ul, i, lds_f64(ua + (unsigned)(C::AT + i) * 8u), lds_f64(ua + (unsigned)(C::BX + 1 + i) * 8u), row_pos(cntm, i));
      }
#endif
      // I(E) at the lane's edge, then the rebinned bin (:847): (I(E_g+1) - I(E_g)) / (E_g+1 - E_g)
      const double I = isTot ? o.S : __fma_rn(o.R, __dsub_rn(E, o.W), o.S);
      double Iprev = __shfl_up_sync(0xffffffffu, I, 1, P);
      if (ul == 0) Iprev = 0.0;
      const double na = triv ? cf : __dmul_rn(__dsub_rn(I, Iprev), idw);
      __syncwarp();  // every lane is done reading a
      if (live) sts_f64(ua + (unsigned)(C::AT + ul) * 8u, na);
      __syncwarp();
    }

    // ---- total tau, w0 clamp, band tau of the unit (:856-886) -> HBM (slot-indexed).  The
    // g-independent terms come from continua_kernel (optics_rorr.cu).
    if (live) {
      const double tau = __dadd_rn(__dadd_rn(q_s3, lds_f64(ua + (unsigned)(C::AT + ul) * 8u)), q_tauc);
      double w0 = 0.0;
      if (!(tau <= RT_TAU_MIN)) w0 = fmin(RT_MAX_W0, q_sc / tau);
      sts_f64(ua + (unsigned)(C::TT + ul) * 8u, tau);
      if (active) {
        ch.tau[obase * G + ul] = tau;
        ch.w0[obase * G + ul] = w0;
      }
    }
    __syncwarp();
    if (ul == 0 && active) {
      double tbnd = 0.0;
#pragma unroll 4
      for (int gg = 0; gg < G; gg++) tbnd = __dadd_rn(tbnd, __dmul_rn(ut[C::TT + gg], s_w[gg]));
      ch.tband[obase] = tbnd;
    }
    if (!unit_ok && active && ul == 0) {
      // unit id in the order the sorting kernels decode (rorr_impl.cuh)
      const unsigned uid = (tb.pad0 & 4) ? (unsigned)obase
                                         : ((((unsigned)c >> 2) * (unsigned)nw + (unsigned)l) * (unsigned)cap + (unsigned)us) * 4u + ((unsigned)c & 3u);
      const int p = atomicAdd(wcount, 1);
      if (p < wcap) worklist[p] = (int)uid;
      st_fail++;
    }
    __syncwarp();
  }
  if (ch.cut_stats) {
    st_fail = __reduce_add_sync(0xffffffffu, st_fail);
    if (lane == 0) {
      atomicAdd(ch.cut_stats + 0, (unsigned long long)st_merges);
      atomicAdd(ch.cut_stats + 1, (unsigned long long)st_passes);
      if (st_restarts) atomicAdd(ch.cut_stats + 2, (unsigned long long)st_restarts);
      if (st_fail) atomicAdd(ch.cut_stats + 3, (unsigned long long)st_fail);
      atomicAdd(ch.cut_stats + 4, (unsigned long long)st_cold);
      if (st_closed) atomicAdd(ch.cut_stats + 6, (unsigned long long)st_closed);
    }
  }
}

template <int P, bool FULL>
void launch(const DevTables &tb, const DevChunk &ch, int mode, int Lc, int *worklist, int *wcount, int wcap,
            cudaStream_t st) {
  using C = Cfg<P>;
  auto kern = rorr_cut_kernel<P, FULL>;
  const size_t smem = (size_t)C::UPB * C::UNIT * sizeof(double) +
                      (size_t)CUT_WPB * (tb.nk - 1) * C::CWN * 32 * sizeof(unsigned);
  static bool attr_set[64] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  if (smem > 48 * 1024 && (dev < 0 || dev >= 64 || !attr_set[dev])) {
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) == cudaSuccess && dev >= 0 && dev < 64)
      attr_set[dev] = true;
  }
  long long chains;
  if (mode == 0) {
    chains = (long long)ch.ncol * ((tb.nw + C::UPW - 1) / C::UPW * C::UPW);
  } else {
    const long long segs = (ch.ncol + Lc - 1) / Lc;
    chains = segs * tb.nw * ((ch.nslot_cap + C::UPW - 1) / C::UPW * C::UPW);
  }
  const long long nb = (chains + C::UPB - 1) / C::UPB;
  if (nb <= 0) return;
  kern<<<(unsigned)nb, 32 * CUT_WPB, smem, st>>>(tb, ch, mode, Lc, worklist, wcount, wcap);
}

}  // namespace rorr_cut

// mode 0: chains over the layer slots of a (column, bin); mode 1: chains over Lc consecutive columns.
// Units that do not settle are appended to the worklist (decoded by the sorting kernels).
void rorr_launch_cut(const DevTables &tb, const DevChunk &ch, int mode, int Lc, int *worklist, int *wcount,
                     int wcap, cudaStream_t st) {
  const int G = tb.G;
  using namespace rorr_cut;
  if (G <= 8) {
    if (G == 8) launch<8, true>(tb, ch, mode, Lc, worklist, wcount, wcap, st);
    else launch<8, false>(tb, ch, mode, Lc, worklist, wcount, wcap, st);
  } else if (G <= 16) {
    if (G == 16) launch<16, true>(tb, ch, mode, Lc, worklist, wcount, wcap, st);
    else launch<16, false>(tb, ch, mode, Lc, worklist, wcount, wcap, st);
  } else {
    if (G == 32) launch<32, true>(tb, ch, mode, Lc, worklist, wcount, wcap, st);
    else launch<32, false>(tb, ch, mode, Lc, worklist, wcount, wcap, st);
  }
}
