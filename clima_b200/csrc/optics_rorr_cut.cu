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
// outside one) and re-evaluated.  The staircases of a unit are the initial guess of the NEXT unit
// of a CHAIN: the same (bin, layer slot) of the next column when neighbouring columns are alike
// (a sweep: 99 % of the warp passes are the only pass), else the next layer slot of the same
// (column, bin) (2.3 passes per merge on randomised columns; experiments/rorr_cut_proto.py).
// The first unit of a chain starts from a 13-step bisection on the value of the edge element.
// A unit that does not settle within the pass budget is appended to the worklist of the exact
// sorting kernel (optics_rorr_exact.cu).
//
// What it costs: 4 shared-memory gathers + ~20 instructions per (edge, row) and pass, 15 x 16 pairs
// per merge, against ~94 instructions per ELEMENT (256 per merge) of the sorting kernel.
#include <cstdlib>

#include "rt_device.cuh"
#include "rt_kernels.h"

#ifndef CUT_WPB
#define CUT_WPB 4  // warps per block
#endif
#ifndef CUT_MIN_BLOCKS
#define CUT_MIN_BLOCKS 6  // 80 registers, 24 warps per SM
#endif
#ifndef CUT_ROW_UNROLL
#define CUT_ROW_UNROLL 1  // count words (4 rows) per trip of the pass loop: fully unrolled it needs 168 registers
#endif

namespace rorr_cut {

template <int P>
struct Cfg {
  static constexpr int LOGP = (P == 8 ? 3 : (P == 16 ? 4 : 5));
  static constexpr int UPW = 32 / P;        // units per warp
  static constexpr int UPB = CUT_WPB * UPW;  // chains per block
  static constexpr int CWN = P / 4;          // count words (4 rows each) per lane and merge
  static constexpr int ROWU = CUT_ROW_UNROLL;
  // per-unit tables, in doubles
  static constexpr int BX = 0;               // [0] = -inf, [1..G] = b, [G+1..2P+1] = +inf
  static constexpr int SB = 2 * P + 2;       // Sb[0..G]
  static constexpr int WB = SB + P + 1;      // Wb[0..G]  (static copy per unit: one gather address)
  static constexpr int AT = WB + P + 1;      // a[0..P-1]
  static constexpr int TT = AT + P;          // scratch: total tau of the unit (band sum)
  static constexpr int UNIT = TT + P;        // doubles per unit
  static constexpr int NBIS = 13;            // bisection steps of a cold start
  static constexpr int MAXPASS = 48;         // passes before a restart from bisection / giving up
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
  int iL, tagR;  // tagR = row | (count of that row << 8)
};

// One pass over the rows for the lane's staircase (count words at cnt, stride 128 bytes).
template <int P, bool FULL>
__device__ __forceinline__ PassOut cut_pass(const unsigned ua, const unsigned cnt, const double (&w)[32], const int G) {
  using C = Cfg<P>;
  PassOut o;
  o.L = __longlong_as_double(0xfff0000000000000ll);
  o.R = __longlong_as_double(0x7ff0000000000000ll);
  o.W = 0.0;
  o.S = 0.0;
  o.iL = 0;
  o.tagR = 0;
#pragma unroll(C::ROWU)
  for (int q = 0; q < C::CWN; q++) {
    const unsigned cw = lds_u32(cnt + q * 128u);
#pragma unroll
    for (int r = 0; r < 4; r++) {
      const int i = 4 * q + r;
      if (FULL || i < G) {
        const unsigned c = (cw >> (8 * r)) & 0xffu;
        const unsigned ad = ua + c * 8u;
        const double bm = lds_f64(ad + C::BX * 8u), bp = lds_f64(ad + C::BX * 8u + 8u);
        const double sb = lds_f64(ad + C::SB * 8u), wb = lds_f64(ad + C::WB * 8u);
        const double ai = lds_f64(ua + (unsigned)(C::AT + i) * 8u);
        const double lo = __dadd_rn(ai, bm), hi = __dadd_rn(ai, bp);  // :828 tau_xy of the two boundary elements
        if (lo > o.L) { o.L = lo; o.iL = i; }
        if (hi < o.R) { o.R = hi; o.tagR = i | (int)(c << 8); }
        const double t1 = __fma_rn(ai, wb, sb);
        o.S = __fma_rn(w[i], t1, o.S);  // w[i]: constant-bank operand (DevTables::wq)
        o.W = __fma_rn(w[i], wb, o.W);
      }
    }
  }
  return o;
}

// number of j with b_j < x (bx[k] = b_(k-1), +inf beyond G)
template <int P>
__device__ __forceinline__ unsigned count_lt(const unsigned ua, const double x) {
  unsigned c = 0;
#pragma unroll
  for (int s = P; s >= 1; s >>= 1)
    if (lds_f64(ua + (Cfg<P>::BX + c + s) * 8u) < x) c += s;
  return c;
}

// Cold start: bisection (on the IEEE-754 pattern) for the value of the element that contains edge
// E, then the staircase of everything below it.  Only an initial guess: cut_pass decides.
// Out of line and rolled: it runs once per chain (and merge), the pass loop must stay small.
template <int P>
__device__ __noinline__ void cut_bisect(const unsigned ua, const unsigned cnt, const unsigned swa,
                                        const int G, const double E) {
  using C = Cfg<P>;
  long long lo = __double_as_longlong(__dadd_rn(lds_f64(ua + C::AT * 8u), lds_f64(ua + (C::BX + 1) * 8u)));
  long long hi = __double_as_longlong(__dadd_rn(lds_f64(ua + (unsigned)(C::AT + G - 1) * 8u), lds_f64(ua + (unsigned)(C::BX + G) * 8u)));
#pragma unroll 1
  for (int it = 0; it < C::NBIS; it++) {
    const double t = __longlong_as_double((lo + hi) >> 1);
    double W = 0.0;
#pragma unroll 4
    for (int i = 0; i < G; i++) {
      const unsigned c = count_lt<P>(ua, __dsub_rn(t, lds_f64(ua + (unsigned)(C::AT + i) * 8u)));
      W = __fma_rn(lds_f64(swa + (unsigned)i * 8u), lds_f64(ua + (C::WB + c) * 8u), W);
    }
    if (W > E) hi = __double_as_longlong(t); else lo = __double_as_longlong(t);
  }
  const double t = __longlong_as_double(lo);
#pragma unroll 1
  for (int q = 0; q < C::CWN; q++) {
    unsigned cw = 0;
#pragma unroll
    for (int r = 0; r < 4; r++) {
      const int i = 4 * q + r;
      if (i < G)
        cw |= count_lt<P>(ua, __dsub_rn(t, lds_f64(ua + (unsigned)(C::AT + i) * 8u))) << (8 * r);
    }
    sts_u32(cnt + q * 128u, cw);
  }
}

// k(g) * column amount of k-gas jj for one unit: bilinear log10k(P,T) interpolation, lane per
// g-point (A.2, :649-662 + interp_2d, linear_interpolation_module.F90:322-335).  Split in two so
// that the table loads are in flight while the previous merge runs.
struct KRaw {
  double f0, f1, f2, f3, q1, q2, col;
};
__device__ __forceinline__ KRaw kgas_load(const DevTables &tb, const DevChunk &ch, const int c, const int l,
                                          const int slot, const int k, const int jj, const int g) {
  const int G = tb.G, nz = ch.nz;
  const int nbr = rt_nbr(tb);
  const int *ib = ch.ibr + ((size_t)c * nz + slot) * nbr;
  const double *qb = ch.qbr + ((size_t)c * nz + slot) * nbr;
  const DevK &kt = tb.k[jj];
  const int ix = ib[2 * jj], iy = ib[2 * jj + 1];
  KRaw r;
  r.q1 = qb[2 * jj];
  r.q2 = qb[2 * jj + 1];
  r.col = ch.dens[(size_t)c * nz * tb.ng + k + (size_t)(kt.sp - 1) * nz] * ch.dz[(size_t)c * nz + k];  // cols(j,sp) :607
  const double *f = kt.log10k + (size_t)G * (ix + (size_t)kt.np * (iy + (size_t)kt.nt * l));
  const size_t sy = (size_t)G * kt.np;
  const int gc = g < G ? g : G - 1;
  r.f0 = f[gc];
  r.f1 = f[gc + G];
  r.f2 = f[gc + sy];
  r.f3 = f[gc + sy + G];
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
  for (int i = ul; i < C::UNIT; i += P) {
    double v = 0.0;
    if (i < C::SB) v = (i == 0) ? __longlong_as_double(0xfff0000000000000ll) : __longlong_as_double(0x7ff0000000000000ll);
    else if (i >= C::WB && i < C::AT) v = tb.wbin_e[(i - C::WB) < G ? (i - C::WB) : G];
    ut[i] = v;
  }
  {
    const unsigned full = (unsigned)G * 0x01010101u;
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
    const bool cold = (n == 0);
    bool unit_ok = true;

    KRaw raw = kgas_load(tb, ch, c, l, us, k, 0, ul);
    for (int jj = 0; jj < nk; jj++) {
      const double v = kgas_finish(raw, live);
      if (jj + 1 < nk) raw = kgas_load(tb, ch, c, l, us, k, jj + 1, ul);
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
        sts_f64(ua + (unsigned)(C::BX + 1 + ul) * 8u, v);
        sts_f64(ua + (unsigned)(C::SB + 1 + ul) * 8u, s);
      }
      __syncwarp();
      const unsigned cntm = cnt0 + (unsigned)((jj - 1) * C::CWN) * 128u;
      if (cold && isEdge) cut_bisect<P>(ua, cntm, swa, G, E);
      // ---- evaluate / repair the staircases
      PassOut o;
      bool ok;
      int passes = 0;
      bool restarted = false;
      while (true) {
        o = cut_pass<P, FULL>(ua, cntm, tb.wq, G);
        const int iR = o.tagR & 0xff, cR = o.tagR >> 8;
        const double we = __dmul_rn(s_w[iR], s_w[cR < G ? cR : G - 1]);  // weight of the element that would come next
        const bool inv = o.L > o.R;
        const bool over = o.W > E + TOL;
        const bool under = E >= __dadd_rn(o.W, we) + TOL;
        ok = !isEdge || !(inv || over || under);
        if (__all_sync(0xffffffffu, ok)) break;
        if (++passes >= C::MAXPASS) {
          if (restarted) break;
          restarted = true;
          passes = 0;
          if (!ok) cut_bisect<P>(ua, cntm, swa, G, E);
          continue;
        }
        if (!ok) {
          if (inv || over) {  // drop the largest element inside
            const unsigned a = cntm + (unsigned)(o.iL >> 2) * 128u;
            sts_u32(a, lds_u32(a) - (1u << (8 * (o.iL & 3))));
          }
          if (inv || (!over && under)) {  // take the smallest element outside
            const unsigned a = cntm + (unsigned)(iR >> 2) * 128u;
            sts_u32(a, lds_u32(a) + (1u << (8 * (iR & 3))));
          }
        }
      }
      {
        const unsigned bal = __ballot_sync(0xffffffffu, ok);
        unit_ok = unit_ok && ((bal & umask) == umask);
      }
      // I(E) at the lane's edge, then the rebinned bin (:847): (I(E_g+1) - I(E_g)) / (E_g+1 - E_g)
      const double I = isTot ? o.S : __fma_rn(o.R, __dsub_rn(E, o.W), o.S);
      double Iprev = __shfl_up_sync(0xffffffffu, I, 1, P);
      if (ul == 0) Iprev = 0.0;
      const double na = __dmul_rn(__dsub_rn(I, Iprev), idw);
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
    }
    __syncwarp();
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
