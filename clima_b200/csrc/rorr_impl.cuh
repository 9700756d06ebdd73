// rorr_impl.cuh -- optical-property assembly + random overlap with resorting and rebinning.
//
// Replaces the body of the per-bin loop of OpticalProperties_compute_opacity
// (src/radtran/clima_radtran_types.f90:638-771) and k_rorr (:780-888) for a batch of columns.
// Included by optics_rorr_fast.cu and optics_rorr_exact.cu, which instantiate the two flavours.
//
// Work unit = (column, wavelength bin l, unique layer slot).  A unit is owned by LPU = P*P/32
// lanes of one warp (P = ngauss rounded up to a power of two: 8 lanes at ngauss=16, 4 units per
// warp).  Per unit:
//   1. continua (Rayleigh, CIA, photolysis xs, H2O continuum, custom, particles)  -- A.3
//      the Rayleigh/CIA/xs/continuum items are dealt round-robin to the unit's lanes and the two
//      partial sums are combined with a butterfly (deterministic, same value on every lane);
//   2. for each k-gas: bilinear log10k(P,T) interpolation, lane per g-point       -- A.2
//   3. RORR merge of the running mixture a(1..G) with the new gas b(1..G): the G x G sums
//      a_i + b_j live in REGISTERS, 32 per lane, and are ranked by a bitonic MERGE network (the G
//      rows are already ascending because k-distributions ascend in g -- checked at table load --
//      and the mixture ascends after every rebin).
//        FAST : 32-bit keys [quantised tau_xy | element index], compare-exchange = 2 x VIMNMX.U32;
//               the rebin walk re-forms every exact tau_xy and certifies that the sequence it walks
//               is non-decreasing; a unit that fails is appended to a worklist.
//        EXACT: 64-bit keys = FP64 bit pattern with the low 2*log2(P) mantissa bits replaced by the
//               element index: one DSETP orders (value, index), i.e. the STABLE rank of mrgrnk.
//   4. weights_to_bins + rebin (clima_eqns.f90:43-54, futils rebin): each lane walks its 32 sorted
//      elements with the running cumulative-weight edge and a monotone output-bin pointer.
//      The gathers a(i), b(j) by sorted element index go to per-warp tables laid out
//      [row][lane]: every lane owns a private COLUMN (bank = lane), so a gather with 32 different
//      rows is conflict-free.  Weights w_i*w_j come from a 4-way replicated wxy table.
//      Per-lane partial bin sums go to shared memory and are added in lane order (no atomics:
//      bit-reproducible run to run).
//   5. total tau, w0 clamp, band tau (:856-886).
#pragma once
#include "rt_device.cuh"
#include "rt_kernels.h"

#ifndef RORR_KPL16
#define RORR_KPL16 32  // keys per lane at P = 16 (64 -> 4 lanes per unit, 3 cross-lane steps instead of 6: measured slower, 168 registers + spills, r1p)
#endif
#ifndef RORR_MIN_BLOCKS
#define RORR_MIN_BLOCKS (RORR_KPL16 == 64 ? 3 : 4)
#endif

#ifndef RORR_WPB
#define RORR_WPB 4  // warps per block; measured r1q per 1024 columns: 2 -> 39.1 ms, 4 -> 34.7 ms, 8 -> 36.1 ms
#endif
#ifndef RORR_NS
#define RORR_NS rorr
#endif

namespace RORR_NS {

template <int P>
struct Cfg {
  static constexpr int LOGP = (P == 8 ? 3 : (P == 16 ? 4 : 5));
  static constexpr int N = P * P;
  static constexpr int KPL = (P == 16 ? RORR_KPL16 : 32);  // keys per lane
  static constexpr int LOGK = (KPL == 32 ? 5 : 6);
  static constexpr int LPU = N / KPL;   // lanes per unit: 2, 8, 32
  static constexpr int GPL = P / LPU;   // g-points per lane: 4, 2, 1
  static constexpr int UPW = 32 / LPU;  // units per warp: 16, 4, 1
  static constexpr int UPB = RORR_WPB * UPW;  // units per block
  static constexpr int NC = (LPU == 32 ? 16 : LPU);  // table columns of one unit
  static constexpr int CW = (LPU == 32 ? 16 : 32);   // table columns per warp
  static constexpr int TBL = P * CW;                 // doubles per table per warp
  static constexpr int SPU = P + 2;                  // edge-integral doubles per unit (G+1, padded)
  static constexpr int WXC = 4;                      // copies of the wxy table
  static constexpr int IDXB = (N <= 256 ? 1 : 2);    // bytes per sorted-index entry
  static constexpr unsigned IDXM = (unsigned)(N - 1);
  // gather addressing: a table is TBLB bytes, aligned to TBLB, row stride 2^ROWSH bytes, so that
  // the byte address of row `row`, column `col` is  base | (row << ROWSH) | (col << 3)  (one LOP3)
  static constexpr int ROWSH = (CW == 32 ? 8 : 7);
  static constexpr unsigned TBLB = (unsigned)TBL * 8u;
  static constexpr unsigned ROWMASK = (unsigned)(P - 1) << ROWSH;
  // dynamic shared memory: [alignment slack TBLB][4 warps x (tA, tB)][wxy N*WXC doubles]
  //                        [4 warps x edge integrals][4 warps x sorted-index bytes (general walk)]
  static constexpr size_t SI_BYTES = (size_t)UPW * 2 * SPU * sizeof(double);  // per unit: I(E_g), then its copy of the edges
  static constexpr size_t IP_BYTES = (size_t)KPL * 32 * IDXB;
  static constexpr size_t BLOCK_BYTES =
      (size_t)TBLB * (2 * RORR_WPB + 1) + (size_t)N * WXC * sizeof(double) + RORR_WPB * (SI_BYTES + IP_BYTES);
};

__device__ __forceinline__ unsigned smem_u32(const void *p) {
  return (unsigned)__cvta_generic_to_shared(p);
}
// shared-memory load by 32-bit address.  volatile: the tables are rewritten between merges, the
// statement must never be merged with an earlier load of the same address.
// Key window: f(v) = ((bits(v) >> S) - (bits(vlo) >> S)) >> s2, a monotone map of [vlo, vhi] onto
// `nbits` bits (S <= 32: one funnel shift of the IEEE-754 pattern; s2 > 0 only when the interval
// spans more than 2^(32+nbits) patterns).
struct KeyWin {
  int S, s2;
  unsigned wb;
};
__device__ __forceinline__ KeyWin key_window(const double vlo, const double vhi, const int nbits) {
  const unsigned long long base = (unsigned long long)__double_as_longlong(vlo);
  const unsigned long long top = (unsigned long long)__double_as_longlong(vhi);
  const int nb = 64 - __clzll((long long)(top - base));  // bit length (0 for an empty range)
  const int shift = nb > nbits ? nb - nbits : 0;
  KeyWin k;
  k.S = shift > 32 ? 32 : shift;
  k.s2 = shift - k.S;
  k.wb = __funnelshift_rc((unsigned)base, (unsigned)(base >> 32), k.S);
  return k;
}
__device__ __forceinline__ unsigned key_of(const double v, const KeyWin &k) {
  return (__funnelshift_rc((unsigned)__double2loint(v), (unsigned)__double2hiint(v), k.S) - k.wb) >> k.s2;
}

template <int B>
__device__ __forceinline__ unsigned lds_idx(unsigned addr) {
  unsigned v;
  if (B == 1) asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(addr));
  else asm volatile("ld.shared.u16 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ void sts_f64(unsigned addr, double v) {
  asm volatile("st.shared.f64 [%0], %1;" ::"r"(addr), "d"(v) : "memory");
}
__device__ __forceinline__ double lds_f64(unsigned addr) {
  double v;
  asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(addr));
  return v;
}

template <int B> struct IdxT { typedef unsigned char type; };
template <> struct IdxT<2> { typedef unsigned short type; };

__device__ __forceinline__ double shfl_xor_d(double v, int lanemask) {
  return __shfl_xor_sync(0xffffffffu, v, lanemask);
}

// compare-exchange, ascending (64-bit keys are distinct by construction)
__device__ __forceinline__ void ce(double &a, double &b) {
  const bool p = a < b;
  const double lo = p ? a : b;
  const double hi = p ? b : a;
  a = lo;
  b = hi;
}
__device__ __forceinline__ void ce(unsigned &a, unsigned &b) {
  const unsigned lo = min(a, b), hi = max(a, b);  // 2 x VIMNMX.U32
  a = lo;
  b = hi;
}
__device__ __forceinline__ double pick(bool keepmin, double a, double y) {
  return ((a < y) == keepmin) ? a : y;
}
__device__ __forceinline__ unsigned pick(bool keepmin, unsigned a, unsigned y) {
  return keepmin ? min(a, y) : max(a, y);
}
__device__ __forceinline__ double shfl_x(double v, int m) { return shfl_xor_d(v, m); }
__device__ __forceinline__ unsigned shfl_x(unsigned v, int m) {
  return __shfl_xor_sync(0xffffffffu, v, m);
}

// Bitonic sort network over positions pos = ul*32 + r, all runs ascending ("flip" variant: the
// first step of stage k pairs pos with pos ^ (2^k-1), the remaining steps pos with pos ^ d).
// Stages below `kfirst` are skipped (rows pre-sorted).  Stages whose partners all live in the lane
// (k <= 5) are unrolled; the cross-lane stages run as a ROLLED loop over k with the lane masks in
// registers, so the code holds one shuffle flip block, one shuffle half-cleaner block and one
// 5-step in-register block instead of a copy per stage (instruction-cache footprint).
template <int P, typename K>
__device__ __forceinline__ void bitonic_network(K (&x)[Cfg<P>::KPL], const int ul, const int kfirst) {
  using C = Cfg<P>;
  constexpr int K1 = 2 * C::LOGP, KPL = C::KPL, LOGK = C::LOGK;
#pragma unroll
  for (int k = 1; k <= (K1 < LOGK ? K1 : LOGK); k++) {
    if (k < kfirst) continue;
    const int m = (1 << k) - 1;
#pragma unroll
    for (int r = 0; r < KPL; r++)
      if ((r & (1 << (k - 1))) == 0) ce(x[r], x[r ^ m]);
#pragma unroll
    for (int s = k - 2; s >= 0; s--) {
#pragma unroll
      for (int r = 0; r < KPL; r++)
        if ((r & (1 << s)) == 0) ce(x[r], x[r | (1 << s)]);
    }
  }
#ifdef RORR_UNROLL_CROSS
#pragma unroll
#else
#pragma unroll 1
#endif
  for (int k = (kfirst > LOGK + 1 ? kfirst : LOGK + 1); k <= K1; k++) {
    {  // flip step: pos <-> pos ^ (2^k - 1): partner lane ul ^ lm, partner register r ^ (KPL-1)
      const int lm = ((1 << k) - 1) >> LOGK;
      const bool keepmin = (ul & ((lm + 1) >> 1)) == 0;
#pragma unroll
      for (int r = 0; r < KPL / 2; r++) {
        const K y1 = shfl_x(x[r ^ (KPL - 1)], lm);
        const K y2 = shfl_x(x[r], lm);
        x[r] = pick(keepmin, x[r], y1);
        x[r ^ (KPL - 1)] = pick(keepmin, x[r ^ (KPL - 1)], y2);
      }
    }
#ifdef RORR_UNROLL_CROSS
#pragma unroll
#else
#pragma unroll 1
#endif
    for (int s = k - 2; s >= LOGK; s--) {  // cross-lane half-cleaners
      const int lm = 1 << (s - LOGK);
      const bool keepmin = (ul & lm) == 0;
#pragma unroll
      for (int r = 0; r < KPL; r++) {
        const K y = shfl_x(x[r], lm);
        x[r] = pick(keepmin, x[r], y);
      }
    }
#pragma unroll
    for (int s = LOGK - 1; s >= 0; s--) {  // in-lane half-cleaners d = KPL/2..1
#pragma unroll
      for (int r = 0; r < KPL; r++)
        if ((r & (1 << s)) == 0) ce(x[r], x[r | (1 << s)]);
    }
  }
}

// Writes value v of table row g into every column of the unit.  Lane `ul` starts at column ul and
// rotates, so the lanes of a half-warp always hit distinct banks.
template <int P>
__device__ __forceinline__ void tbl_put(double *t, const int g, const double v, const int colbase,
                                        const int ul) {
  using C = Cfg<P>;
#pragma unroll
  for (int cc = 0; cc < C::NC; cc++) t[g * C::CW + colbase + ((cc + ul) & (C::NC - 1))] = v;
}

// k(g) * column of k-gas jj for the unit's layer: bilinear log10k(P,T) interpolation, lane per
// g-point (A.2, :649-662 + interp_2d, linear_interpolation_module.F90:322-335).  Split in two so
// that the table loads of the NEXT gas are in flight while the current merge builds its keys:
//   kgas_load   issues the loads (bracket record, column amount, 4 table corners per g-point)
//   kgas_finish interpolates, exponentiates and scales; the caller puts the values into a table.
template <int P>
struct KgasRaw {
  double f[4 * Cfg<P>::GPL];
  double q1, q2, col;
};

template <int P>
__device__ __forceinline__ void kgas_load(const DevTables &tb, const DevChunk &ch, const int c,
                                          const int l, const int slot, const int k, const int jj,
                                          const int ul, KgasRaw<P> &r) {
  using C = Cfg<P>;
  const int G = tb.G, nz = ch.nz;
  const int nbr = rt_nbr(tb);
  const int *ib = ch.ibr + ((size_t)c * nz + slot) * nbr;
  const double *qb = ch.qbr + ((size_t)c * nz + slot) * nbr;
  const DevK &kt = tb.k[jj];
  const int ix = ib[2 * jj], iy = ib[2 * jj + 1];
  r.q1 = qb[2 * jj];
  r.q2 = qb[2 * jj + 1];
  r.col = ch.dens[(size_t)c * nz * tb.ng + k + (size_t)(kt.sp - 1) * nz] *
          ch.dz[(size_t)c * nz + k];  // cols(j,sp) :607
  const double *f = kt.log10k + (size_t)G * (ix + (size_t)kt.np * (iy + (size_t)kt.nt * l));
  const size_t sy = (size_t)G * kt.np;
#pragma unroll
  for (int t = 0; t < C::GPL; t++) {
    const int g = ul + C::LPU * t;
    const int gc = g < G ? g : G - 1;
    r.f[4 * t + 0] = f[gc];
    r.f[4 * t + 1] = f[gc + G];
    r.f[4 * t + 2] = f[gc + sy];
    r.f[4 * t + 3] = f[gc + sy + G];
  }
}

template <int P>
__device__ __forceinline__ void kgas_finish(const KgasRaw<P> &r, const int G, const int ul,
                                            double (&v)[Cfg<P>::GPL]) {
  using C = Cfg<P>;
  const double p1 = 1.0 - r.q1, p2 = 1.0 - r.q2;
#pragma unroll
  for (int t = 0; t < C::GPL; t++) {
    const int g = ul + C::LPU * t;
    const double fx1 = p1 * r.f[4 * t + 0] + r.q1 * r.f[4 * t + 1];
    const double fx2 = p1 * r.f[4 * t + 2] + r.q1 * r.f[4 * t + 3];
    const double lk = p2 * fx1 + r.q2 * fx2;
    // lean exp inside the physical range of k-tables, the library routine otherwise
    const double kv = (lk > -280.0 && lk < 280.0) ? rt_ten2power_lean(lk) : rt_ten2power_cold(lk);
    v[t] = g < G ? kv * r.col : 0.0;  // :660, :816-819 / :828
  }
}

// weights_to_bins + rebin (:843-849) of the unit's sorted sequence.  ip[r*32] (this lane's column
// of the warp's index buffer) holds the element index (i*P + j) of sorted position ul*32 + r; the
// lanes of the unit own consecutive runs of 32 positions; `tot` is the lane's weight total.
//   scan  : exclusive scan of the lane totals = cumulative weight at the lane's first element
//           (weights_to_bins);
//   walk  : the 32 elements with the running cumulative weight o and the running integral
//           I = sum w*tau.  The rebinned value of new bin g is (I(E_g+1) - I(E_g)) / (E_g+1 - E_g)
//           (futils rebin: overlap-weighted mean), so the walk only has to record I at the G-1
//           interior edges: the element that contains edge E is split there,
//           I(E) = I_before + (E - o_before) * tau.  Every edge lies in exactly one element of one
//           lane (lane l owns (off_l, off_l+1]), hence one writer per edge, no partial sums, no
//           atomics.  I is lane-relative during the walk; the lane offsets come from a second scan.
//           The gathers of the next element are issued before the current one is processed.
// The new bin values are written to table A (all columns of the unit).
// CERT: also certify that the exact tau_xy sequence is non-decreasing (fast kernel); returns it.
template <int P, bool FULL, bool CERT>
__device__ __forceinline__ bool rt_rebin(const typename IdxT<Cfg<P>::IDXB>::type *__restrict__ ip,
                                         double *tA, const double *tB, double *sI,
                                         const double *wx, const double *s_E, const double *s_ib,
                                         const double tot, const int G, const int ul,
                                         const int mycol, const int colbase) {
  using C = Cfg<P>;
  constexpr int LPU = C::LPU, LOGP = C::LOGP, CW = C::CW;
  double off = tot;
#pragma unroll
  for (int d = 1; d < LPU; d <<= 1) {
    const double t = __shfl_up_sync(0xffffffffu, off, d, LPU);
    if (ul >= d) off = __dadd_rn(off, t);
  }
  // lane l owns the edges in (off_l, off_l+1]: its last element ends exactly at the next lane's start
  // (bit-exactly: off of lane l+1 IS off_next of lane l, otherwise an edge within rounding of a
  // lane boundary -- common: E_8 = 0.5 = the weight of the first 8 rows -- gets two owners or none)
  const double off_next = off;
  off = __shfl_up_sync(0xffffffffu, off_next, 1, LPU);
  if (ul == 0) off = 0.0;
  int g = 0;  // largest g with E[g] < off  (E[g+1] >= off)
#pragma unroll
  for (int s = P / 2; s > 0; s >>= 1)
    if (g + s <= G - 1 && s_E[g + s] < off) g += s;
  const int g_first = g;
  double n1 = s_E[g + 1];
  double I = 0.0, o0 = off, prev = -1.0, first = -1.0;
  bool ok = true;
  const double *cA = tA + mycol;
  const double *cB = tB + mycol;
  unsigned idx = ip[0];
  double na = cA[(idx >> LOGP) * CW], nb = cB[(idx & (P - 1)) * CW], nw = wx[idx * C::WXC];
#pragma unroll 4
  for (int r = 0; r < C::KPL; r++) {
    const double a = na, b = nb, w = nw;
    if (r < C::KPL - 1) {
      idx = ip[(r + 1) * 32];
      na = cA[(idx >> LOGP) * CW];
      nb = cB[(idx & (P - 1)) * CW];
      nw = wx[idx * C::WXC];
    }
    const double tv = __dadd_rn(a, b);  // exact tau_xy (:828); padding elements: w = 0
    if (CERT) {
      if (FULL || w > 0.0) {
        ok = ok && (tv >= prev);
        prev = tv;
        if (r == 0) first = tv;
      }
    }
    const double o1 = (r == C::KPL - 1 && ul < LPU - 1) ? off_next : __dadd_rn(o0, w);
    while (o1 > n1 && g < G - 1) {  // the element contains the right edge of new bin g
      g++;
      sI[g] = __fma_rn(__dsub_rn(n1, o0), tv, I);
      n1 = s_E[g + 1];
    }
    I = __fma_rn(w, tv, I);
    o0 = o1;
  }
  // lane offsets of the integral
  double Ioff = I;
#pragma unroll
  for (int d = 1; d < LPU; d <<= 1) {
    const double t = __shfl_up_sync(0xffffffffu, Ioff, d, LPU);
    if (ul >= d) Ioff = __dadd_rn(Ioff, t);
  }
  if (ul == LPU - 1) sI[G] = Ioff;  // I(E_G) = the whole integral
  if (ul == 0) sI[0] = 0.0;
  Ioff = __shfl_up_sync(0xffffffffu, Ioff, 1, LPU);
  if (ul == 0) Ioff = 0.0;
  for (int e = g_first + 1; e <= g; e++) sI[e] = __dadd_rn(sI[e], Ioff);
  if (CERT) {  // lane boundary: this lane's first element against the previous lane's last one
    const double plast = __shfl_up_sync(0xffffffffu, prev, 1, LPU);
    if (ul > 0 && first >= 0.0) ok = ok && (first >= plast);
  }
  __syncwarp();  // edge integrals complete; every lane is done reading tA
#pragma unroll
  for (int t = 0; t < C::GPL; t++) {
    const int gg = ul + LPU * t;
    double v = 0.0;
    if (gg < G) v = __dmul_rn(__dsub_rn(sI[gg + 1], sI[gg]), s_ib[gg]);
    tbl_put<P>(tA, gg, v, colbase, ul);
  }
  __syncwarp();
  return ok;
}

// Same contract as rt_rebin, specialised for weight sets in which one sorted element can contain
// AT MOST ONE bin edge (max w_i*w_j < min wbin: true for every Gauss-Legendre set; checked on the
// host at table load, DevTables::pad0 bit 1 selects the general walk otherwise).  Differences:
//   * the single possible edge of an element is an if-converted block (no inner loop, no divergent
//     branch: with 32 lanes walking, SOME lane has an edge in 85 % of the steps);
//   * gathers by 32-bit shared address  base | row offset  (tables aligned to their size), issued
//     one element ahead of the recurrence;
//   * every unit has its own copy of the bin edges behind its I(E_g) slots, with +inf beyond the
//     last interior edge (no g < G-1 test); the edge block walks ONE pointer through both;
//   * a unit with `triv` set receives `cf` (its closed-form result) instead of the rebinned values.
// Rolled (x4): the kernel is instruction-cache bound, every unrolled copy costs more than it saves.
template <int P, bool FULL, bool CERT>
__device__ __forceinline__ bool rt_rebin1(const unsigned aIP, double *tA, double *sI,
                                          const unsigned aA, const unsigned aB, const unsigned aW,
                                          const double *s_ib, const double tot,
                                          const int G, const int ul, const int colbase,
                                          const bool triv, const double (&cf)[Cfg<P>::GPL]) {
  using C = Cfg<P>;
  constexpr int LPU = C::LPU, LOGP = C::LOGP, KPL = C::KPL, ROWSH = C::ROWSH;
  double off = tot;
#pragma unroll
  for (int d = 1; d < LPU; d <<= 1) {
    const double t = __shfl_up_sync(0xffffffffu, off, d, LPU);
    if (ul >= d) off = __dadd_rn(off, t);
  }
  const double off_next = off;  // bit-identically the start of the next lane (edge ownership)
  off = __shfl_up_sync(0xffffffffu, off_next, 1, LPU);
  if (ul == 0) off = 0.0;
  const double *sEn = sI + C::SPU;  // the unit's copy of the edges (+inf beyond the last interior one)
  int g = 0;  // largest g with E[g] < off
#pragma unroll
  for (int s = P / 2; s > 0; s >>= 1)
    if (sEn[g + s] < off) g += s;
  const int g_first = g;
  // pg -> I(E_g) of the current bin (32-bit shared address); the edge block is
  //   pg[1] = I(edge); n1 = next edge; pg++
  unsigned pg = smem_u32(sI + g);
  double n1 = lds_f64(pg + (unsigned)(C::SPU + 1) * 8u);
  double I = 0.0, o0 = off, prev = -1.0;
  bool ok = true;
  unsigned idx = lds_idx<C::IDXB>(aIP);
  double na = lds_f64(aA | ((idx << (ROWSH - LOGP)) & C::ROWMASK));
  double nb = lds_f64(aB | ((idx << ROWSH) & C::ROWMASK));
  double nw = lds_f64(aW + idx * (unsigned)(C::WXC * 8));
  const double first = __dadd_rn(na, nb);
#pragma unroll 4
  for (int r = 0; r < KPL - 1; r++) {
    const double a = na, b = nb, w = nw;
    idx = lds_idx<C::IDXB>(aIP + (unsigned)(r + 1) * (32u * C::IDXB));
    na = lds_f64(aA | ((idx << (ROWSH - LOGP)) & C::ROWMASK));
    nb = lds_f64(aB | ((idx << ROWSH) & C::ROWMASK));
    nw = lds_f64(aW + idx * (unsigned)(C::WXC * 8));
    const double tv = __dadd_rn(a, b);  // exact tau_xy (:828); padding elements: w = 0
    if (CERT) {
      if (FULL || w > 0.0) {
        ok = ok && (tv >= prev);
        prev = tv;
      }
    }
    const double o1 = __dadd_rn(o0, w);
    if (o1 > n1) {  // the element contains the right edge of new bin g
      sts_f64(pg + 8u, __fma_rn(__dsub_rn(n1, o0), tv, I));
      n1 = lds_f64(pg + (unsigned)(C::SPU + 2) * 8u);
      pg += 8u;
    }
    I = __fma_rn(w, tv, I);
    o0 = o1;
  }
  {  // last element: ends bit-exactly where the next lane starts
    const double tv = __dadd_rn(na, nb);
    if (CERT) {
      if (FULL || nw > 0.0) {
        ok = ok && (tv >= prev);
        prev = tv;
      }
    }
    const double o1 = ul < LPU - 1 ? off_next : __dadd_rn(o0, nw);
    if (o1 > n1) {
      sts_f64(pg + 8u, __fma_rn(__dsub_rn(n1, o0), tv, I));
      pg += 8u;
    }
    I = __fma_rn(nw, tv, I);
  }
  g = (int)((pg - smem_u32(sI)) >> 3);
  // lane offsets of the integral
  double Ioff = I;
#pragma unroll
  for (int d = 1; d < LPU; d <<= 1) {
    const double t = __shfl_up_sync(0xffffffffu, Ioff, d, LPU);
    if (ul >= d) Ioff = __dadd_rn(Ioff, t);
  }
  if (ul == LPU - 1) sI[G] = Ioff;  // I(E_G) = the whole integral
  if (ul == 0) sI[0] = 0.0;
  Ioff = __shfl_up_sync(0xffffffffu, Ioff, 1, LPU);
  if (ul == 0) Ioff = 0.0;
  for (int e = g_first + 1; e <= g; e++) sI[e] = __dadd_rn(sI[e], Ioff);
  if (CERT) {  // lane boundary: this lane's first element against the previous lane's last one
    const double plast = __shfl_up_sync(0xffffffffu, prev, 1, LPU);
    if (ul > 0 && (FULL || prev >= 0.0)) ok = ok && (first >= plast);
  }
  __syncwarp();  // edge integrals complete; every lane is done reading tA
#pragma unroll
  for (int t = 0; t < C::GPL; t++) {
    const int gg = ul + LPU * t;
    double v = 0.0;
    if (gg < G) v = __dmul_rn(__dsub_rn(sI[gg + 1], sI[gg]), s_ib[gg]);
    tbl_put<P>(tA, gg, triv ? cf[t] : v, colbase, ul);
  }
  __syncwarp();
  return ok;
}

// ---------------------------------------------------------------------------------------------
// The kernel.  FAST: 32-bit quantised keys + certificate, failed units appended to `worklist`.
// !FAST: exact 64-bit (value,index) keys; if `worklist` is given, only the listed units.
//   key32 = [ f : 32-2*log2(P) bits ][ element index pos : 2*log2(P) bits ],
//   f     = (bits(tau_xy) - bits(min tau_xy)) >> shift, shift chosen so that f of the largest sum
//           (a_G + b_G; a and b ascend) still fits.
// f is a monotone quantisation of tau_xy (IEEE-754 bit patterns of non-negative doubles order like
// integers).  Quantisation can only mis-order sums whose f collide; elements with equal f sit in
// index order, which is the stable order of `mrgrnk` whenever they are exactly equal or come from
// the same row.  The walk recomputes every EXACT tau_xy and checks the sequence is non-decreasing;
// otherwise the unit is redone by the exact kernel.  The fast path therefore returns the same
// permutation as the reference's stable merge sort or hands the unit over.
// ---------------------------------------------------------------------------------------------
template <int P, bool FULL, bool FAST>
__global__ void __launch_bounds__(32 * RORR_WPB, RORR_MIN_BLOCKS * 4 / RORR_WPB)
optics_rorr_kernel(const __grid_constant__ DevTables tb, const __grid_constant__ DevChunk ch,
                   int *__restrict__ worklist, int *__restrict__ wcount, const int wcap) {
  using C = Cfg<P>;
  typedef typename IdxT<C::IDXB>::type idx_t;
  constexpr int LPU = C::LPU, UPB = C::UPB, LOGP = C::LOGP, CW = C::CW;
  constexpr int IDXBITS = 2 * LOGP, FB = 32 - IDXBITS;
  __shared__ double s_w[P];       // wbin, zero padded
  __shared__ double s_E[P + 1];   // wbin_e
  __shared__ double s_ib[P];      // 1/(wbin_e(g+1)-wbin_e(g))
  extern __shared__ double s_dyn[];

  const int G = FULL ? P : tb.G, nw = tb.nw, nz = ch.nz;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int ul = lane & (LPU - 1);   // lane within unit
  const int ub = threadIdx.x / LPU;  // unit within block
  const int mycol = (LPU == 32) ? (lane & 15) : lane;
  const int colbase = (LPU == 32) ? 0 : (lane - ul);
  // tables aligned to their own size (gather address = base | row offset)
  const unsigned dyn_u = smem_u32(s_dyn);
  const unsigned tb0 = (dyn_u + C::TBLB - 1u) & ~(C::TBLB - 1u);
  char *dbase = (char *)s_dyn + (tb0 - dyn_u);
  double *tA = (double *)(dbase + (size_t)warp * 2 * C::TBLB);  // running mixture a(i), [row][column]
  double *tB = tA + C::TBL;                                      // new gas b(j)
  double *s_wxy = (double *)(dbase + 2 * RORR_WPB * (size_t)C::TBLB);  // [idx][copy]: wxy(j + i*P), types_create.f90:216-222
  double *sI = s_wxy + C::N * C::WXC + ((size_t)warp * C::UPW + (ub % C::UPW)) * (2 * C::SPU);  // I(E_g), then edges
  idx_t *ip = (idx_t *)(s_wxy + C::N * C::WXC + RORR_WPB * (C::UPW * 2 * C::SPU)) + (size_t)warp * (C::KPL * 32) + lane;
  const double *wx = s_wxy + (lane & (C::WXC - 1));
  const unsigned aA = smem_u32(tA + mycol), aB = smem_u32(tB + mycol), aW = smem_u32(wx), aIP = smem_u32(ip);

  for (int i = threadIdx.x; i < P; i += blockDim.x) {
    s_w[i] = i < G ? tb.wbin[i] : 0.0;
    s_ib[i] = i < G ? tb.inv_dw[i] : 0.0;
  }
  for (int i = threadIdx.x; i < P + 1; i += blockDim.x) s_E[i] = tb.wbin_e[i < G ? i : G];
  for (int i = threadIdx.x; i < UPB * (P + 1); i += blockDim.x) {  // per-unit edge copies
    const int u = i / (P + 1), e = i % (P + 1);
    s_wxy[C::N * C::WXC + (size_t)u * (2 * C::SPU) + C::SPU + e] =
        (e == 0) ? 0.0 : (e <= G - 1 ? tb.wbin_e[e] : __longlong_as_double(0x7ff0000000000000ll));
  }
  // wxy(j + i*P) = wbin(i)*wbin(j) (types_create.f90:216-222), WXC copies per entry, precomputed
  // on the host in exactly this layout (capi.cu)
  for (int i = threadIdx.x; i < C::N * C::WXC / 2; i += blockDim.x)
    reinterpret_cast<double2 *>(s_wxy)[i] = reinterpret_cast<const double2 *>(tb.wxy)[i];
  __syncthreads();

  // unit ids fit in 31 bits (the chunk is sized so, capi.cu): 32-bit divisions, no 64-bit subroutines
  unsigned uid = blockIdx.x * (unsigned)UPB + (unsigned)ub;
  const unsigned cap = (unsigned)ch.nslot_cap;
  bool active = true;
  if (!FAST && worklist) {
    int n = *wcount;
    if (n > wcap) n = wcap;
    if (uid < (unsigned)n) uid = (unsigned)worklist[uid]; else { active = false; uid = 0; }
  }
  // unit id -> (column, bin, slot).  Two orders.  Column-grouped: the units of a warp are the SAME
  // (bin, slot) of 4 adjacent columns (uid = ((c/4 * nw + l) * cap + slot) * 4 + c%4): neighbouring
  // columns of a sweep take the closed-form / sort decision together and read the same k-table rows
  // (cfg2: 34.6 vs 35.2 ms per 1024 columns).  Slot-fastest: 4 adjacent slots of ONE column, better
  // when neighbouring columns are unrelated (cfg5, randomised composition: 37.9 vs 45.0 ms).
  // similarity_kernel counts the 4-column groups with unlike k-gas amounts; the host picks
  // column-grouped only if fewer than 1 in 16 are unlike and says so in pad0 bit 2 (slot-fastest).
  int slot, l;
  unsigned cl;
  {
    const bool byslot = (tb.pad0 & 4) != 0;
    const unsigned q = byslot ? uid : (uid >> 2);
    const unsigned q1 = q / cap;
    slot = (int)(q - q1 * cap);
    const unsigned q2 = q1 / (unsigned)nw;
    l = (int)(q1 - q2 * (unsigned)nw);
    cl = byslot ? q2 : q2 * 4u + (uid & 3u);
  }
  if (cl >= (unsigned)ch.ncol) active = false;
  const int c = cl < (unsigned)ch.ncol ? (int)cl : ch.ncol - 1;
  const int nsl = ch.nslots[c];
  if (slot >= nsl) {
    active = false;
    slot = nsl - 1;
  }
  if (__all_sync(0xffffffffu, !active)) return;

  const int k = ch.layer_of_slot[(size_t)c * nz + slot];  // ground-first layer index
  const unsigned umask = (LPU == 32 ? 0xffffffffu : ((1u << LPU) - 1u)) << (lane & ~(LPU - 1));
  bool unit_ok = true;
  const int kfirst = (FAST || tb.rows_sorted) ? LOGP + 1 : 1;
  const double *cA = tA + mycol, *cB = tB + mycol;
  const bool general_walk = (tb.pad0 & 2) != 0;

  // Gas pipeline: at the top of iteration jj `bn` holds k(g)*column of gas jj; the table loads of gas
  // jj+1 are issued there and finished (interpolation + exp) at ONE place in the loop body.
  // Iteration 0 only installs gas 0 as the running mixture.
  double bn[C::GPL];
  {
    KgasRaw<P> r0;
    kgas_load<P>(tb, ch, c, l, slot, k, 0, ul, r0);
    kgas_finish<P>(r0, G, ul, bn);
  }
  for (int jj = 0; jj < tb.nk; jj++) {
    {
      double *tdst = jj == 0 ? tA : tB;
#pragma unroll
      for (int t = 0; t < C::GPL; t++) tbl_put<P>(tdst, ul + LPU * t, bn[t], colbase, ul);
    }
    __syncwarp();
    KgasRaw<P> rn;
    const bool more = jj + 1 < tb.nk;
    if (more) kgas_load<P>(tb, ch, c, l, slot, k, jj + 1, ul, rn);

    // ---- merges that need no sort.  If the rows do not interleave (a_i + b_G <= a_i+1 + b_1 for
    // every i) the stable sorted order IS the row-major index order, row i covers exactly the
    // cumulative-weight interval of new bin i, and the rebin (:843-849) returns
    // a_i * sum_j(w_j) + sum_j(w_j b_j) = a_i + <b>; symmetrically b_j + <a> when the columns do not
    // interleave.  (The reference's result differs from the closed form only by the rounding of
    // its 256-term cumulative weights, O(1e-15) relative.)  The warp skips the sort as a whole when
    // every unit of the warp qualifies.
    bool triv = false, urow = false;
    double cf[C::GPL];
    if (jj > 0) {
      const double a1 = cA[0], aG = cA[(G - 1) * CW], b1 = cB[0], bG = cB[(G - 1) * CW];
      bool rowmaj = true, colmaj = true;
      double ag[C::GPL], bg[C::GPL];
#pragma unroll
      for (int t = 0; t < C::GPL; t++) {
        const int g = ul + LPU * t;
        const int gn = g + 1 < P ? g + 1 : g;
        ag[t] = cA[g * CW];
        bg[t] = cB[g * CW];
        if (g < G - 1) {
          rowmaj = rowmaj && (ag[t] + bG <= cA[gn * CW] + b1);
          colmaj = colmaj && (bg[t] + aG <= cB[gn * CW] + a1);
        }
      }
      const unsigned br = __ballot_sync(0xffffffffu, rowmaj), bc = __ballot_sync(0xffffffffu, colmaj);
      urow = (br & umask) == umask;
      triv = !(tb.pad0 & 1) && (urow || (bc & umask) == umask);
      // A qualifying unit ALWAYS takes the closed form, also when its warp sorts because of another
      // unit: the result of a unit must not depend on its warp-mates (the exact kernel regroups units).
      if (__any_sync(0xffffffffu, triv)) {
        double mean = 0.0;  // <b> or <a>, summed in g order
        const double *cm = urow ? cB : cA;
        for (int g = 0; g < G; g++) mean = __fma_rn(s_w[g], cm[g * CW], mean);
#pragma unroll
        for (int t = 0; t < C::GPL; t++)
          cf[t] = (ul + LPU * t) < G ? __dadd_rn(urow ? ag[t] : bg[t], mean) : 0.0;
      }
    }
    if (more) kgas_finish<P>(rn, G, ul, bn);
    if (jj == 0) continue;
    if (__all_sync(0xffffffffu, triv)) {
      __syncwarp();
#pragma unroll
      for (int t = 0; t < C::GPL; t++) tbl_put<P>(tA, ul + LPU * t, cf[t], colbase, ul);
      __syncwarp();
      continue;
    }

    // ---- keys: pos = ul*KPL + r ;  i = pos >> LOGP (running mixture), j = pos & (P-1) (new gas)
    double tot = 0.0;
    if (FAST) {
      // One key window over [a_1+b_1, a_G+b_G] resolves ~2^-20 relative.  That is not enough where
      // the TOP rows of the mixture form a plateau (saturated k-distributions: a_p..a_G equal to
      // 1e-4 relative) and the new gas is decades weaker: the sums of those rows differ only in
      // their excess over A = a_p + b_1.  Such a unit gets a second window over the EXCESS
      // (v - A) + d, d = A*2^-46 (regime bit | 23-bit window; experiments/rorr_stats.py: units that
      // fail the certificate 8.9 % -> 1.6 % on the cfg2 workload).  The warp takes the two-window
      // key loop only if one of its units has a plateau.
      const double vmin = cA[0] + cB[0], vmax = cA[(G - 1) * CW] + cB[(G - 1) * CW];
      int prow = 0;
      {
        const double thr = cA[(G - 1) * CW] * (1.0 - 0x1p-12);
        int hi = G - 1;
        while (prow < hi) {  // first row with a >= thr (rows ascend)
          const int mid = (prow + hi) >> 1;
          if (cA[mid * CW] < thr) prow = mid + 1; else hi = mid;
        }
      }
      const bool two = prow < G - 1;
      unsigned x[C::KPL];
      double bj[P <= 16 ? P : 1], ai[P <= 16 ? C::KPL / P : 1];  // the lane's operands, loaded once
      if (P <= 16) {
#pragma unroll
        for (int j = 0; j < P; j++) bj[j] = cB[j * CW];
#pragma unroll
        for (int t = 0; t < C::KPL / P; t++) ai[t] = cA[(ul * (C::KPL / P) + t) * CW];
      }
      if (!__any_sync(0xffffffffu, two)) {
        const KeyWin kw = key_window(vmin, vmax, FB);
#pragma unroll
        for (int r = 0; r < C::KPL; r++) {
          const int pos = ul * C::KPL + r;
          const int i = pos >> LOGP, j = pos & (P - 1);
          const double v = (P <= 16) ? __dadd_rn(ai[P <= 16 ? r / P : 0], bj[P <= 16 ? r % P : 0])
                                     : __dadd_rn(cA[i * CW], cB[j * CW]);  // :828 tau_xy
          const unsigned key = (key_of(v, kw) << IDXBITS) | (unsigned)pos;
          x[r] = (FULL || ((i < G) && (j < G))) ? key : 0xffffffffu;
        }
      } else {
        const double Apiv = two ? __dadd_rn(cA[prow * CW], cB[0]) : __longlong_as_double(0x7ff0000000000000ll);
        const double dlt = two ? Apiv * 0x1p-46 : 0.0;
        const KeyWin kl = key_window(vmin, two ? Apiv : vmax, two ? FB - 1 : FB);
        const KeyWin ku = key_window(dlt, two ? __dadd_rn(__dsub_rn(vmax, Apiv), dlt) : dlt, FB - 1);
#pragma unroll
        for (int r = 0; r < C::KPL; r++) {
          const int pos = ul * C::KPL + r;
          const int i = pos >> LOGP, j = pos & (P - 1);
          const double v = (P <= 16) ? __dadd_rn(ai[P <= 16 ? r / P : 0], bj[P <= 16 ? r % P : 0])
                                     : __dadd_rn(cA[i * CW], cB[j * CW]);  // :828 tau_xy
          const double e = __dadd_rn(__dsub_rn(v, Apiv), dlt);
          const unsigned fl = key_of(v, kl), fu = key_of(e, ku) | (1u << (FB - 1));
          const unsigned key = ((v >= Apiv ? fu : fl) << IDXBITS) | (unsigned)pos;
          x[r] = (FULL || ((i < G) && (j < G))) ? key : 0xffffffffu;
        }
      }
      bitonic_network<P, unsigned>(x, ul, kfirst);
#pragma unroll
      for (int r = 0; r < C::KPL; r++) {
        const unsigned idx = x[r] & C::IDXM;
        ip[r * 32] = (idx_t)idx;
        tot = __dadd_rn(tot, lds_f64(aW + idx * (unsigned)(C::WXC * 8)));
      }
    } else {
      double x[C::KPL];
      double bj[P <= 16 ? P : 1], ai[P <= 16 ? C::KPL / P : 1];  // the lane's operands, loaded once
      if (P <= 16) {
#pragma unroll
        for (int j = 0; j < P; j++) bj[j] = cB[j * CW];
#pragma unroll
        for (int t = 0; t < C::KPL / P; t++) ai[t] = cA[(ul * (C::KPL / P) + t) * CW];
      }
#pragma unroll
      for (int r = 0; r < C::KPL; r++) {
        const int pos = ul * C::KPL + r;
        const int i = pos >> LOGP, j = pos & (P - 1);
        const double v = (P <= 16) ? __dadd_rn(ai[P <= 16 ? r / P : 0], bj[P <= 16 ? r % P : 0])
                                   : __dadd_rn(cA[i * CW], cB[j * CW]);  // :828 tau_xy
        const unsigned long long bits = (FULL || ((i < G) && (j < G)))
                                            ? (unsigned long long)__double_as_longlong(v)
                                            : 0x7fe0000000000000ull;  // huge, finite
        x[r] = __longlong_as_double(
            (long long)((bits & ~(unsigned long long)C::IDXM) | (unsigned long long)pos));
      }
      bitonic_network<P, double>(x, ul, kfirst);
#pragma unroll
      for (int r = 0; r < C::KPL; r++) {
        const unsigned idx = (unsigned)__double2loint(x[r]) & C::IDXM;
        ip[r * 32] = (idx_t)idx;
        tot = __dadd_rn(tot, lds_f64(aW + idx * (unsigned)(C::WXC * 8)));
      }
    }
    __syncwarp();
    bool ok;
    if (!general_walk) {
      ok = rt_rebin1<P, FULL, FAST>(aIP, tA, sI, aA, aB, aW, s_ib, tot, G, ul, colbase, triv, cf);
    } else {
      ok = rt_rebin<P, FULL, FAST>(ip, tA, tB, sI, wx, s_E, s_ib, tot, G, ul, mycol, colbase);
      if (__any_sync(0xffffffffu, triv)) {  // closed form for the qualifying units of a sorting warp
        if (triv) {
#pragma unroll
          for (int t = 0; t < C::GPL; t++) tbl_put<P>(tA, ul + LPU * t, cf[t], colbase, ul);
        }
        __syncwarp();
      }
    }
    if (FAST) {
      const unsigned bal = __ballot_sync(0xffffffffu, ok || triv);
      unit_ok = unit_ok && ((bal & umask) == umask);
    }
  }

  // ---- total tau, w0 clamp, band tau of the unit (:856-886) -> HBM (slot-indexed).  The
  // g-independent terms come from continua_kernel (optics_rorr.cu).
  const size_t obase = (((size_t)c * tb.nw + l) * ch.nslot_cap + slot);
  const size_t U = (size_t)ch.ncol * tb.nw * ch.nslot_cap;
  const double q_s3 = ch.cont[obase], q_tauc = ch.cont[U + obase], q_sc = ch.cont[2 * U + obase];
#pragma unroll
  for (int t = 0; t < C::GPL; t++) {
    const int gg = ul + LPU * t;
    if (gg < G) {
      const double tau = __dadd_rn(__dadd_rn(q_s3, cA[gg * CW]), q_tauc);
      double w0 = 0.0;
      if (!(tau <= RT_TAU_MIN)) w0 = fmin(RT_MAX_W0, q_sc / tau);
      sI[gg] = tau;
      if (active) {
        ch.tau[obase * G + gg] = tau;
        ch.w0[obase * G + gg] = w0;
      }
    }
  }
  __syncwarp();
  if (ul == 0 && active) {
    double tbnd = 0.0;
#pragma unroll 4
    for (int gg = 0; gg < G; gg++) tbnd = __dadd_rn(tbnd, __dmul_rn(sI[gg], s_w[gg]));
    ch.tband[obase] = tbnd;
  }
  if (FAST && !unit_ok && active && ul == 0) {
    const int p = atomicAdd(wcount, 1);
    if (p < wcap) worklist[p] = (int)uid;
  }
}

}  // namespace RORR_NS
