// optics_rorr.cu -- optical-property assembly + random overlap with resorting and rebinning.
//
// Replaces the body of the per-bin loop of OpticalProperties_compute_opacity
// (src/radtran/clima_radtran_types.f90:638-771) and k_rorr (:780-888) for a batch of columns.
//
// Work unit = (column, wavelength bin l, unique layer slot).  A unit is owned by LPU = P*P/32
// lanes of one warp (P = ngauss rounded up to a power of two: 8 lanes at ngauss=16, 4 units per
// warp).  Per unit:
//   1. continua (Rayleigh, CIA, photolysis xs, H2O continuum, custom, particles)  -- A.3
//   2. for each k-gas: bilinear log10k(P,T) interpolation, lane per g-point       -- A.2
//   3. RORR merge of the running mixture a(1..G) with the new gas b(1..G):
//        keys(i,j) = a_i + b_j.  The G x G keys live in REGISTERS, 32 per lane, as FP64 bit
//        patterns whose low 2*log2(P) mantissa bits are replaced by the element index (i,j):
//        one 64-bit compare then orders (value, index) lexicographically, i.e. a STABLE rank like
//        mrgrnk (futils), needs no payload, and perturbs only the ORDER of keys closer than
//        2^-44 relative -- the exact tau is re-formed from (i,j) after the sort.
//        Rows are already ascending because k-distributions ascend in g (checked at table load,
//        `rows_sorted`), so the bitonic network starts at merge stage log2(P)+1: 26 of the 36
//        compare-exchange steps at P=16, 20 of them in-register, 6 through warp shuffles.
//   4. weights_to_bins + rebin (clima_eqns.f90:43-54, futils rebin): per-lane running edge over
//        its 32 sorted elements, monotone output-bin pointer, deterministic per-lane partials in
//        shared memory (no atomics: bit-reproducible run to run).
//   5. total tau, w0 clamp, band tau (:856-886).
#include "rt_device.cuh"
#include "rt_kernels.h"

namespace {

template <int P>
struct Cfg {
  static constexpr int LOGP = (P == 8 ? 3 : (P == 16 ? 4 : 5));
  static constexpr int N = P * P;
  static constexpr int KPL = 32;       // keys per lane
  static constexpr int LPU = N / KPL;  // lanes per unit: 2, 8, 32
  static constexpr int GPL = P / LPU;  // g-points per lane in the interpolation: 4, 2, 1
  static constexpr int UPB = 128 / LPU;  // units per 128-thread block
  static constexpr unsigned long long IDXMASK = (unsigned long long)(N - 1);
  static constexpr int ISTR = N + LPU;  // u16 stride of a unit's sorted-index list (bank spread)
};

__device__ __forceinline__ double shfl_xor_d(double v, int lanemask) {
  return __shfl_xor_sync(0xffffffffu, v, lanemask);
}

// compare-exchange, ascending (keys are distinct by construction)
__device__ __forceinline__ void ce(double &a, double &b) {
  const bool p = a < b;
  const double lo = p ? a : b;
  const double hi = p ? b : a;
  a = lo;
  b = hi;
}

// Bitonic sort network over positions pos = ul*32 + r, all runs ascending ("flip" variant: the first
// step of stage k pairs pos with pos ^ (2^k-1), the remaining steps pos with pos ^ d).  Stages
// below `kfirst` are skipped (rows pre-sorted).
template <int P>
__device__ __forceinline__ void bitonic_network(double (&x)[32], const int ul, const int kfirst) {
  using C = Cfg<P>;
#pragma unroll
  for (int k = 1; k <= 2 * C::LOGP; k++) {
    if (k < kfirst) continue;
    // ---- flip step
    if (k <= 5) {
      const int m = (1 << k) - 1;
#pragma unroll
      for (int r = 0; r < 32; r++)
        if ((r & (1 << (k - 1))) == 0) ce(x[r], x[r ^ m]);
    } else {
      const int lm = ((1 << k) - 1) >> 5;
      const bool keepmin = (ul & ((lm + 1) >> 1)) == 0;
#pragma unroll
      for (int r = 0; r < 16; r++) {
        const double y1 = shfl_xor_d(x[r ^ 31], lm);
        const double y2 = shfl_xor_d(x[r], lm);
        x[r] = ((x[r] < y1) == keepmin) ? x[r] : y1;
        x[r ^ 31] = ((x[r ^ 31] < y2) == keepmin) ? x[r ^ 31] : y2;
      }
    }
    // ---- half-cleaner steps
#pragma unroll
    for (int s = k - 2; s >= 0; s--) {
      const int d = 1 << s;
      if (d < 32) {
#pragma unroll
        for (int r = 0; r < 32; r++)
          if ((r & d) == 0) ce(x[r], x[r | d]);
      } else {
        const int lm = d >> 5;
        const bool keepmin = (ul & lm) == 0;
#pragma unroll
        for (int r = 0; r < 32; r++) {
          const double y = shfl_xor_d(x[r], lm);
          x[r] = ((x[r] < y) == keepmin) ? x[r] : y;
        }
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Pieces shared by the fast and the exact kernel.
// ---------------------------------------------------------------------------------------------
struct Continua {
  double tausg, taua, taup, tausp, tauc, tausc, gt;
};

// Rayleigh, CIA, photolysis, water continuum, particles, custom opacity and the asymmetry
// parameter of one (column c, bin l, slot) -- A.3, clima_radtran_types.f90:664-757.
__device__ __forceinline__ Continua rt_continua(const DevTables &tb, const DevChunk &ch, const int c,
                                                const int l, const int slot, const int k) {
  const int nw = tb.nw, nz = ch.nz, ng = tb.ng;
  const double *dens = ch.dens + (size_t)c * nz * ng + k;  // dens[i*nz]
  const double dzk = ch.dz[(size_t)c * nz + k];
  const int nbr = rt_nbr(tb);
  const int *ib = ch.ibr + ((size_t)c * nz + slot) * nbr;
  const double *qb = ch.qbr + ((size_t)c * nz + slot) * nbr;
  int rbase = 2 * tb.nk;
  Continua o;
  // Rayleigh (:686-693)
  double tausg = 0.0;
#pragma unroll 1
  for (int i = 0; i < tb.nray; i++) {
    const double col = dens[(size_t)(tb.ray_sp[i] - 1) * nz] * dzk;
    tausg = tausg + tb.ray_xs[l + (size_t)i * nw] * col;
  }
  // CIA (:696-704), photolysis (:707-713)
  double taua = 0.0;
#pragma unroll 1
  for (int i = 0; i < tb.ncia + tb.npxs; i++) {
    const bool is_cia = i < tb.ncia;
    const DevXS &x = is_cia ? tb.cia[i] : tb.pxs[i - tb.ncia];
    double xs;
    if (x.dim == 0) {
      xs = x.data[l];
    } else {
      const int it = ib[rbase + i];
      const double q1 = qb[rbase + i], p1 = 1.0 - q1;
      const double *f = x.data + (size_t)l * x.nt;
      xs = rt_ten2power(p1 * f[it] + q1 * f[it + 1]);
    }
    if (is_cia)
      taua = taua + xs * dens[(size_t)(x.sp1 - 1) * nz] * dens[(size_t)(x.sp2 - 1) * nz] * dzk;
    else
      taua = taua + xs * (dens[(size_t)(x.sp1 - 1) * nz] * dzk);
  }
  rbase += tb.ncia + tb.npxs;
  // water continuum (:716-723)
  if (tb.has_cont) {
    const int it = ib[rbase];
    const double q1 = qb[rbase], p1 = 1.0 - q1;
    const double *fh = tb.cont_H2O + (size_t)l * tb.cont_nt;
    const double *ff = tb.cont_foreign + (size_t)l * tb.cont_nt;
    const double H2O = rt_ten2power(p1 * fh[it] + q1 * fh[it + 1]);
    const double foreign = rt_ten2power(p1 * ff[it] + q1 * ff[it + 1]);
    const double dw = dens[(size_t)(tb.cont_LH2O - 1) * nz];
    taua = taua + H2O * dw * (dw * dzk) + foreign * dw * ch.foreign_col[(size_t)c * nz + slot];
  }
  rbase += 1;
  // particles (:733-744)
  double tausp = 0.0, taup = 0.0;
  double tausp_1[RT_MAXPART], gtp[RT_MAXPART];
  if (tb.npart > 0) {
    const double *rad = ch.radii + (size_t)c * nz * tb.np + k;
    const double *pdn = ch.pdens + (size_t)c * nz * tb.np + k;
#pragma unroll 1
    for (int i = 0; i < tb.npart; i++) {
      const DevPart &p = tb.part[i];
      const int ir = ib[rbase + i];
      const double q1 = qb[rbase + i], p1 = 1.0 - q1;
      const size_t o2 = (size_t)l * p.nr + ir;
      const double w0p = p1 * p.w0[o2] + q1 * p.w0[o2 + 1];
      const double qext = p1 * p.qext[o2] + q1 * p.qext[o2 + 1];
      gtp[i] = p1 * p.gt[o2] + q1 * p.gt[o2 + 1];
      const double r = rad[(size_t)(p.p_ind - 1) * nz];
      const double taup_1 = qext * RT_PI * (r * r) * pdn[(size_t)(p.p_ind - 1) * nz] * dzk;
      taup = taup + taup_1;
      tausp_1[i] = w0p * taup_1;
      tausp = tausp + tausp_1[i];
    }
  }
  rbase += tb.npart;
  // custom optical properties (:726-730, 546-572)
  double tauc = RT_TINY, w0c = RT_TINY, g0c = RT_TINY;
  if (tb.has_custom) {
    const int ip = ib[rbase];
    const double q1 = qb[rbase], p1 = 1.0 - q1;
    const size_t o2 = (size_t)l * tb.custom_nP + ip;
    tauc = (p1 * tb.c_dtau[o2] + q1 * tb.c_dtau[o2 + 1]) * dzk;
    w0c = p1 * tb.c_w0[o2] + q1 * tb.c_w0[o2 + 1];
    g0c = p1 * tb.c_g0[o2] + q1 * tb.c_g0[o2 + 1];
  }
  const double tausc = w0c * tauc;
  // asymmetry (:746-757)
  double gt = 0.0;
  {
    const double den = fmax(RT_TAU_MIN, (tausp + tausg + tausc));
#pragma unroll 1
    for (int i = 0; i < tb.npart; i++) gt = gt + gtp[i] * tausp_1[i] / den;
    gt = gt + g0c * tausc / den;
    gt = fmin(gt, RT_MAX_GT);
  }
  o.tausg = tausg; o.taua = taua; o.taup = taup; o.tausp = tausp; o.tauc = tauc; o.tausc = tausc;
  o.gt = gt;
  return o;
}

// k(g) * column of k-gas jj for the unit's layer: bilinear log10k(P,T) interpolation, lane per
// g-point (A.2, :649-662 + interp_2d, linear_interpolation_module.F90:322-335), into dst[0..P).
template <int P>
__device__ __forceinline__ void rt_kgas(const DevTables &tb, const DevChunk &ch, const int c,
                                        const int l, const int slot, const int k, const int jj,
                                        const int ul, double *dst) {
  using C = Cfg<P>;
  const int G = tb.G, nz = ch.nz;
  const int nbr = rt_nbr(tb);
  const int *ib = ch.ibr + ((size_t)c * nz + slot) * nbr;
  const double *qb = ch.qbr + ((size_t)c * nz + slot) * nbr;
  const DevK &kt = tb.k[jj];
  const int ix = ib[2 * jj], iy = ib[2 * jj + 1];
  const double q1 = qb[2 * jj], q2 = qb[2 * jj + 1];
  const double p1 = 1.0 - q1, p2 = 1.0 - q2;
  const double col = ch.dens[(size_t)c * nz * tb.ng + k + (size_t)(kt.sp - 1) * nz] *
                     ch.dz[(size_t)c * nz + k];  // cols(j,sp) :607
  const double *f = kt.log10k + (size_t)G * (ix + (size_t)kt.np * (iy + (size_t)kt.nt * l));
  const size_t sy = (size_t)G * kt.np;
#pragma unroll
  for (int t = 0; t < C::GPL; t++) {
    const int g = ul + C::LPU * t;
    if (g < G) {
      const double fx1 = p1 * f[g] + q1 * f[g + G];
      const double fx2 = p1 * f[g + sy] + q1 * f[g + sy + G];
      const double v = p2 * fx1 + q2 * fx2;
      dst[g] = rt_ten2power(v) * col;  // :660, :816-819 / :828
    } else {
      dst[g] = 0.0;
    }
  }
}

// weights_to_bins + rebin (:843-849) of the unit's sorted sequence.  `sidx[r * LPU + ul]` holds the
// element index (i*P + j) of sorted position ul*32 + r; the lanes of the unit own consecutive runs
// of 32 positions.  Compact on purpose (rolled loops): the unrolled version of this pass made the
// kernel 3x larger than the instruction cache and "no instruction" its top stall.
//   pass 1: lane weight totals -> exclusive scan over the unit's lanes = cumulative weight at the
//           lane's first element (weights_to_bins);
//   pass 2: walk the 32 elements with the running edge; an element inside the current new bin adds
//           w*tau, one that crosses a new-bin edge is split at the edge (futils rebin); every closed
//           bin is scaled by 1/width and written to the lane's partial slot.
// The new bin values land in sa[0..P).  Deterministic: per-lane partials, summed in lane order.
// CERT: also certify that the exact tau_xy sequence is non-decreasing (fast kernel); returns it.
template <int P, bool CERT>
__device__ __forceinline__ bool rt_rebin(const unsigned short *__restrict__ sidx, double *sa,
                                         const double *sb, double *sp, const double *s_wxy,
                                         const double *s_E, const double *s_ib, const int G,
                                         const int ul) {
  using C = Cfg<P>;
  constexpr int LPU = C::LPU, LOGP = C::LOGP;
  double tot = 0.0;
#pragma unroll 8
  for (int r = 0; r < 32; r++) tot = __dadd_rn(tot, s_wxy[sidx[r * LPU + ul]]);
  double off = tot;
#pragma unroll
  for (int d = 1; d < LPU; d <<= 1) {
    const double t = __shfl_up_sync(0xffffffffu, off, d, LPU);
    if (ul >= d) off = __dadd_rn(off, t);
  }
  off = __dsub_rn(off, tot);
  for (int g = 0; g < G; g++) sp[g * LPU + ul] = 0.0;
  int g = 0;
#pragma unroll
  for (int s = P / 2; s > 0; s >>= 1)
    if (g + s <= G - 1 && s_E[g + s] <= off) g += s;
  double n1 = s_E[g + 1];
  double acc = 0.0, o0 = off, prev = -1.0, first = -1.0;
  bool ok = true;
#pragma unroll 2
  for (int r = 0; r < 32; r++) {
    const unsigned idx = sidx[r * LPU + ul];
    const double w = s_wxy[idx];
    if (w > 0.0) {  // padding elements (ngauss < P) carry zero weight
      const double tv = __dadd_rn(sa[idx >> LOGP], sb[idx & (P - 1)]);  // exact tau_xy (:828)
      if (CERT) {
        ok = ok && (tv >= prev);
        prev = tv;
        if (r == 0) first = tv;
      }
      const double o1 = __dadd_rn(o0, w);
      if (o1 > n1 && g < G - 1) {
        double lo = o0;
        do {  // the element crosses the right edge of new bin g
          acc = __fma_rn(__dsub_rn(n1, lo), tv, acc);
          sp[g * LPU + ul] = __dmul_rn(acc, s_ib[g]);
          acc = 0.0;
          g++;
          lo = n1;
          n1 = s_E[g + 1];
        } while (o1 > n1 && g < G - 1);
        acc = __fma_rn(__dsub_rn(o1, lo), tv, acc);
      } else {
        acc = __fma_rn(w, tv, acc);
      }
      o0 = o1;
    }
  }
  sp[g * LPU + ul] = __dmul_rn(acc, s_ib[g]);
  if (CERT) {  // lane boundary: this lane's first element against the previous lane's last one
    const double plast = __shfl_up_sync(0xffffffffu, prev, 1, LPU);
    if (ul > 0 && first >= 0.0) ok = ok && (first >= plast);
  }
  __syncwarp();
#pragma unroll
  for (int t = 0; t < C::GPL; t++) {
    const int gg = ul + LPU * t;
    double sum = 0.0;
    if (gg < G) {
#pragma unroll
      for (int q = 0; q < LPU; q++) sum = __dadd_rn(sum, sp[gg * LPU + q]);
    }
    sa[gg] = sum;
  }
  __syncwarp();
  return ok;
}

// total tau, w0 clamp, band tau, asymmetry of the unit (:856-886) -> HBM (slot-indexed layout).
template <int P>
__device__ __forceinline__ void rt_store_unit(const DevTables &tb, const DevChunk &ch, const Continua &q,
                                              const int c, const int l, const int slot, const int ul,
                                              const bool active, const double *sa, double *sb,
                                              const double *s_w) {
  using C = Cfg<P>;
  const int G = tb.G;
  const size_t obase = (((size_t)c * tb.nw + l) * ch.nslot_cap + slot);
#pragma unroll
  for (int t = 0; t < C::GPL; t++) {
    const int gg = ul + C::LPU * t;
    if (gg < G) {
      const double tau = __dadd_rn(__dadd_rn(__dadd_rn(__dadd_rn(q.tausg, q.taua), q.taup), sa[gg]), q.tauc);
      double w0 = 0.0;
      if (!(tau <= RT_TAU_MIN)) w0 = fmin(RT_MAX_W0, (q.tausg + q.tausp + q.tausc) / tau);
      sb[gg] = tau;
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
    for (int gg = 0; gg < G; gg++) tbnd = __dadd_rn(tbnd, __dmul_rn(sb[gg], s_w[gg]));
    ch.tband[obase] = tbnd;
    ch.gt[obase] = q.gt;
  }
}

// ---------------------------------------------------------------------------------------------
// EXACT kernel: full 64-bit (value, index) bitonic network.  Used for every unit when the tables
// are not ascending in g, and otherwise only for the units the fast kernel could not certify
// (`worklist`, filled by the fast kernel).
// ---------------------------------------------------------------------------------------------
template <int P>
__global__ void __launch_bounds__(128)
optics_rorr_kernel(const __grid_constant__ DevTables tb, const __grid_constant__ DevChunk ch,
                   const int *__restrict__ worklist, const int *__restrict__ wcount, const int wcap) {
  using C = Cfg<P>;
  constexpr int LPU = C::LPU, UPB = C::UPB, LOGP = C::LOGP;
  constexpr int AST = P + 1;              // padded strides: units of one warp hit distinct banks
  constexpr int PST = P * LPU + 1;
  __shared__ double s_w[P];               // wbin, zero padded
  __shared__ double s_E[P + 1];           // wbin_e
  __shared__ double s_ib[P];              // 1/(wbin_e(g+1)-wbin_e(g))
  __shared__ double s_a[UPB * AST];       // running mixture tau_k(g)
  __shared__ double s_b[UPB * AST];       // new gas k*col (g)
  extern __shared__ double s_part[];      // [UPB * PST] rebin partials [g][lane] (dynamic: 33 KB at P=32)
  __shared__ double s_wxy[P * P];         // wxy(j + i*P) = wbin(i)*wbin(j), types_create.f90:216-222
  __shared__ unsigned short s_idx[UPB * C::ISTR];  // sorted element indices [r][lane] per unit

  const int G = tb.G, nw = tb.nw, nz = ch.nz;
  for (int i = threadIdx.x; i < P; i += blockDim.x) {
    s_w[i] = i < G ? tb.wbin[i] : 0.0;
    s_ib[i] = i < G ? tb.inv_dw[i] : 0.0;
  }
  for (int i = threadIdx.x; i < P + 1; i += blockDim.x) s_E[i] = tb.wbin_e[i < G ? i : G];
  for (int i = threadIdx.x; i < P * P; i += blockDim.x) {
    const int a = i >> LOGP, b = i & (P - 1);
    s_wxy[i] = (a < G && b < G) ? __dmul_rn(tb.wbin[a], tb.wbin[b]) : 0.0;
  }
  __syncthreads();

  const int lane = threadIdx.x & 31;
  const int ul = lane & (LPU - 1);     // lane within unit
  const int ub = threadIdx.x / LPU;    // unit within block
  long long uid = (long long)blockIdx.x * UPB + ub;
  const int cap = ch.nslot_cap;
  bool active = true;
  if (worklist) {
    int n = *wcount;
    if (n > wcap) n = wcap;
    if (uid < n) uid = worklist[uid]; else { active = false; uid = 0; }
  }
  int slot = (int)(uid % cap);
  int l = (int)((uid / cap) % nw);
  long long cl = uid / ((long long)cap * nw);
  if (cl >= ch.ncol) active = false;
  int c = cl < ch.ncol ? (int)cl : ch.ncol - 1;
  const int nsl = ch.nslots[c];
  if (slot >= nsl) {
    active = false;
    slot = nsl - 1;
  }
  if (__all_sync(0xffffffffu, !active)) return;

  double *sa = s_a + ub * AST, *sb = s_b + ub * AST, *sp = s_part + ub * PST;
  const int k = ch.layer_of_slot[(size_t)c * nz + slot];  // ground-first layer index
  const Continua q = rt_continua(tb, ch, c, l, slot, k);

  const int kfirst = tb.rows_sorted ? LOGP + 1 : 1;
  for (int jj = 0; jj < tb.nk; jj++) {
    rt_kgas<P>(tb, ch, c, l, slot, k, jj, ul, jj == 0 ? sa : sb);
    __syncwarp();
    if (jj == 0) continue;

    // ---- keys: pos = ul*32 + r ;  i = pos >> LOGP (running mixture), j = pos & (P-1) (new gas)
    double x[32];
#pragma unroll
    for (int r = 0; r < 32; r++) {
      const int pos = ul * 32 + r;
      const int i = pos >> LOGP, j = pos & (P - 1);
      const bool valid = (i < G) && (j < G);
      const double v = sa[i] + sb[j];  // :828 tau_xy
      const unsigned long long bits =
          valid ? (unsigned long long)__double_as_longlong(v) : 0x7fe0000000000000ull;  // huge, finite
      x[r] = __longlong_as_double((long long)((bits & ~C::IDXMASK) | (unsigned long long)pos));
    }
    bitonic_network<P>(x, ul, kfirst);

    unsigned short *sidx = s_idx + ub * C::ISTR;
#pragma unroll
    for (int r = 0; r < 32; r++)
      sidx[r * LPU + ul] =
          (unsigned short)((unsigned long long)__double_as_longlong(x[r]) & C::IDXMASK);
    __syncwarp();
    rt_rebin<P, false>(sidx, sa, sb, sp, s_wxy, s_E, s_ib, G, ul);
  }
  rt_store_unit<P>(tb, ch, q, c, l, slot, ul, active, sa, sb, s_w);
}

// ---------------------------------------------------------------------------------------------
// FAST kernel.  Same work unit and lane mapping, but the G x G sums are ranked with 32-bit keys:
//   key = [ f : 32-2*log2(P) bits ][ element index pos : 2*log2(P) bits ],
//   f   = (bits(tau_xy) - bits(min tau_xy)) >> shift,  shift chosen so that f of the largest sum
//         (a_G + b_G; a and b ascend) still fits.
// f is a monotone, order-preserving quantisation of tau_xy (IEEE-754 bit patterns of non-negative
// doubles order like integers), so one VIMNMX.U32 orders (f, index) lexicographically and a
// compare-exchange costs 2 instructions instead of DSETP + 4 selects, with half the registers.
// Quantisation can only mis-order sums whose f collide; the rebin pass recomputes every EXACT
// tau_xy from its (i,j) anyway and checks that the sequence it walks is non-decreasing.  Elements
// with equal f sit in index order, which is the stable order of `mrgrnk` whenever they are exactly
// equal or come from the same row; if any unit sees a decrease it is appended to `worklist` and
// redone by the exact kernel.  The fast path therefore returns the same permutation as the
// reference's stable merge sort or hands the unit over -- it never silently differs.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void ce_u32(unsigned &a, unsigned &b) {
  const unsigned lo = min(a, b), hi = max(a, b);  // 2 x VIMNMX.U32
  a = lo;
  b = hi;
}

// Merge stages LOGP+1 .. 2*LOGP of the bitonic network ("flip" variant, rows pre-sorted) over
// positions pos = ul*32 + r.  Stages whose partners all live in the lane (k <= 5) are unrolled;
// the cross-lane stages run as a ROLLED loop over k with the lane masks in registers, so the code
// holds one shuffle flip block, one shuffle half-cleaner block and one 5-step in-register block
// instead of a copy per stage (instruction-cache footprint, see rt_rebin).
template <int P>
__device__ __forceinline__ void bitonic_network_u32(unsigned (&x)[32], const int ul) {
  using C = Cfg<P>;
  constexpr int K0 = C::LOGP + 1, K1 = 2 * C::LOGP;
#pragma unroll
  for (int k = K0; k <= (K1 < 5 ? K1 : 5); k++) {
    const int m = (1 << k) - 1;
#pragma unroll
    for (int r = 0; r < 32; r++)
      if ((r & (1 << (k - 1))) == 0) ce_u32(x[r], x[r ^ m]);
#pragma unroll
    for (int s = k - 2; s >= 0; s--) {
#pragma unroll
      for (int r = 0; r < 32; r++)
        if ((r & (1 << s)) == 0) ce_u32(x[r], x[r | (1 << s)]);
    }
  }
#pragma unroll 1
  for (int k = (K0 > 6 ? K0 : 6); k <= K1; k++) {
    {  // flip step: pos <-> pos ^ (2^k - 1): partner lane ul ^ lm, partner register r ^ 31
      const int lm = ((1 << k) - 1) >> 5;
      const bool keepmin = (ul & ((lm + 1) >> 1)) == 0;
#pragma unroll
      for (int r = 0; r < 16; r++) {
        const unsigned y1 = __shfl_xor_sync(0xffffffffu, x[r ^ 31], lm);
        const unsigned y2 = __shfl_xor_sync(0xffffffffu, x[r], lm);
        x[r] = keepmin ? min(x[r], y1) : max(x[r], y1);
        x[r ^ 31] = keepmin ? min(x[r ^ 31], y2) : max(x[r ^ 31], y2);
      }
    }
#pragma unroll 1
    for (int s = k - 2; s >= 5; s--) {  // cross-lane half-cleaners
      const int lm = 1 << (s - 5);
      const bool keepmin = (ul & lm) == 0;
#pragma unroll
      for (int r = 0; r < 32; r++) {
        const unsigned y = __shfl_xor_sync(0xffffffffu, x[r], lm);
        x[r] = keepmin ? min(x[r], y) : max(x[r], y);
      }
    }
#pragma unroll
    for (int s = 4; s >= 0; s--) {  // in-lane half-cleaners d = 16..1
#pragma unroll
      for (int r = 0; r < 32; r++)
        if ((r & (1 << s)) == 0) ce_u32(x[r], x[r | (1 << s)]);
    }
  }
}

template <int P>
__global__ void __launch_bounds__(128, 4)
optics_rorr_fast_kernel(const __grid_constant__ DevTables tb, const __grid_constant__ DevChunk ch,
                        int *__restrict__ worklist, int *__restrict__ wcount, const int wcap) {
  using C = Cfg<P>;
  constexpr int LPU = C::LPU, UPB = C::UPB, LOGP = C::LOGP;
  constexpr int IDXBITS = 2 * LOGP, FB = 32 - IDXBITS;
  constexpr unsigned IDXM = (unsigned)(P * P - 1);
  constexpr int AST = P + 1;
  constexpr int PST = P * LPU + 1;
  __shared__ double s_w[P];
  __shared__ double s_E[P + 1];
  __shared__ double s_ib[P];
  __shared__ double s_wxy[P * P];         // wxy(j + i*P) = wbin(i)*wbin(j), types_create.f90:216-222
  __shared__ double s_a[UPB * AST];
  __shared__ double s_b[UPB * AST];
  extern __shared__ double s_part[];  // [UPB * PST]
  __shared__ unsigned short s_idx[UPB * C::ISTR];

  const int G = tb.G, nw = tb.nw, nz = ch.nz;
  for (int i = threadIdx.x; i < P; i += blockDim.x) {
    s_w[i] = i < G ? tb.wbin[i] : 0.0;
    s_ib[i] = i < G ? tb.inv_dw[i] : 0.0;
  }
  for (int i = threadIdx.x; i < P + 1; i += blockDim.x) s_E[i] = tb.wbin_e[i < G ? i : G];
  for (int i = threadIdx.x; i < P * P; i += blockDim.x) {
    const int a = i >> LOGP, b = i & (P - 1);
    s_wxy[i] = (a < G && b < G) ? __dmul_rn(tb.wbin[a], tb.wbin[b]) : 0.0;
  }
  __syncthreads();

  const int lane = threadIdx.x & 31;
  const int ul = lane & (LPU - 1);
  const int ub = threadIdx.x / LPU;
  const long long uid = (long long)blockIdx.x * UPB + ub;
  const int cap = ch.nslot_cap;
  int slot = (int)(uid % cap);
  int l = (int)((uid / cap) % nw);
  long long cl = uid / ((long long)cap * nw);
  bool active = cl < ch.ncol;
  int c = active ? (int)cl : ch.ncol - 1;
  const int nsl = ch.nslots[c];
  if (slot >= nsl) {
    active = false;
    slot = nsl - 1;
  }
  if (__all_sync(0xffffffffu, !active)) return;

  double *sa = s_a + ub * AST, *sb = s_b + ub * AST, *sp = s_part + ub * PST;
  const int k = ch.layer_of_slot[(size_t)c * nz + slot];
  const Continua q = rt_continua(tb, ch, c, l, slot, k);
  const unsigned umask = (LPU == 32 ? 0xffffffffu : ((1u << LPU) - 1u)) << (lane & ~(LPU - 1));
  bool unit_ok = true;

  for (int jj = 0; jj < tb.nk; jj++) {
    rt_kgas<P>(tb, ch, c, l, slot, k, jj, ul, jj == 0 ? sa : sb);
    __syncwarp();
    if (jj == 0) continue;

    // ---- 32-bit keys
    const unsigned long long base = (unsigned long long)__double_as_longlong(sa[0] + sb[0]);
    const unsigned long long top = (unsigned long long)__double_as_longlong(sa[G - 1] + sb[G - 1]);
    const unsigned long long range = top - base;
    const int nb = 64 - __clzll((long long)range);  // bit length (0 for range == 0)
    const int shift = nb > FB ? nb - FB : 0;
    unsigned x[32];
#pragma unroll
    for (int r = 0; r < 32; r++) {
      const int pos = ul * 32 + r;
      const int i = pos >> LOGP, j = pos & (P - 1);
      const bool valid = (i < G) && (j < G);
      const double v = sa[i] + sb[j];  // :828 tau_xy
      const unsigned f = (unsigned)(((unsigned long long)__double_as_longlong(v) - base) >> shift);
      x[r] = valid ? ((f << IDXBITS) | (unsigned)pos) : 0xffffffffu;
    }
    bitonic_network_u32<P>(x, ul);

    // ---- weights_to_bins + rebin with the monotonicity certificate
    unsigned short *sidx = s_idx + ub * C::ISTR;
#pragma unroll
    for (int r = 0; r < 32; r++) sidx[r * LPU + ul] = (unsigned short)(x[r] & IDXM);
    __syncwarp();
    const bool ok = rt_rebin<P, true>(sidx, sa, sb, sp, s_wxy, s_E, s_ib, G, ul);
    const unsigned bal = __ballot_sync(0xffffffffu, ok);
    unit_ok = unit_ok && ((bal & umask) == umask);
  }
  rt_store_unit<P>(tb, ch, q, c, l, slot, ul, active, sa, sb, s_w);
  if (!unit_ok && active && ul == 0) {
    const int p = atomicAdd(wcount, 1);
    if (p < wcap) worklist[p] = (int)uid;
  }
}

// Expand the slot-indexed opacity of one column into the reference's OpticalPropertiesResult layout
// tau(nz,G,nw), w0(nz,G,nw), g(nz,nw), tau_band(nz,nw), top-first (clima_radtran_types.f90:242-247).
__global__ void expand_opacity_kernel(const DevTables tb, const DevChunk ch, int c, double *tau,
                                      double *w0, double *g, double *tband) {
  const int nz = ch.nz, G = tb.G, nw = tb.nw, cap = ch.nslot_cap;
  const size_t total = (size_t)nz * G * nw;
  for (size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x; t < total;
       t += (size_t)gridDim.x * blockDim.x) {
    const int n = (int)(t % nz);
    const int gg = (int)((t / nz) % G);
    const int l = (int)(t / ((size_t)nz * G));
    const int k = nz - 1 - n;
    const int slot = ch.slot_of_layer[(size_t)c * nz + k];
    const size_t o = (((size_t)c * nw + l) * cap + slot);
    tau[t] = ch.tau[o * G + gg];
    w0[t] = ch.w0[o * G + gg];
    if (gg == 0) {
      g[n + (size_t)l * nz] = ch.gt[o];
      tband[n + (size_t)l * nz] = ch.tband[o];
    }
  }
}

// tau_band of one channel, ground-first (nz,nw_ch,ncol) (clima_radtran_radiate.f90:151-154)
__global__ void tau_band_kernel(const DevTables tb, const DevChunk ch, int ind_start, int nwch,
                                double *out) {
  const int nz = ch.nz, nw = tb.nw, cap = ch.nslot_cap;
  const size_t total = (size_t)nz * nwch * ch.ncol;
  for (size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x; t < total;
       t += (size_t)gridDim.x * blockDim.x) {
    const int i = (int)(t % nz);
    const int ll = (int)((t / nz) % nwch);
    const int c = (int)(t / ((size_t)nz * nwch));
    const int slot = ch.slot_of_layer[(size_t)c * nz + i];
    out[t] = ch.tband[((size_t)c * nw + (ind_start + ll)) * cap + slot];
  }
}

}  // namespace

template <int P>
static void launch_rorr_p(const DevTables &tb, const DevChunk &ch, int *worklist, int *wcount, int wcap,
                          int exact_only, cudaStream_t st) {
  const long long units = (long long)ch.ncol * tb.nw * ch.nslot_cap;
  const long long nb = (units + Cfg<P>::UPB - 1) / Cfg<P>::UPB;
  constexpr size_t dsm = (size_t)Cfg<P>::UPB * (P * Cfg<P>::LPU + 1) * sizeof(double);
  static bool attr_set = false;
  if (!attr_set) {
    cudaFuncSetAttribute(optics_rorr_kernel<P>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dsm);
    cudaFuncSetAttribute(optics_rorr_fast_kernel<P>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dsm);
    attr_set = true;
  }
  if (exact_only || !tb.rows_sorted) {
    optics_rorr_kernel<P><<<(unsigned)nb, 128, dsm, st>>>(tb, ch, nullptr, nullptr, 0);
    return;
  }
  cudaMemsetAsync(wcount, 0, sizeof(int), st);
  optics_rorr_fast_kernel<P><<<(unsigned)nb, 128, dsm, st>>>(tb, ch, worklist, wcount, wcap);
  const long long nbx = ((long long)wcap + Cfg<P>::UPB - 1) / Cfg<P>::UPB;
  optics_rorr_kernel<P><<<(unsigned)nbx, 128, dsm, st>>>(tb, ch, worklist, wcount, wcap);
}

// exact_only != 0: every unit through the 64-bit exact kernel.  Otherwise fast kernel, then the
// exact kernel over the units the fast kernel could not certify (worklist of capacity wcap; the
// caller checks *wcount <= wcap afterwards and re-runs with exact_only on overflow).
void rt_launch_optics_rorr(const DevTables &tb, const DevChunk &ch, int *worklist, int *wcount,
                           int wcap, int exact_only, cudaStream_t st) {
  if (tb.G <= 8) launch_rorr_p<8>(tb, ch, worklist, wcount, wcap, exact_only, st);
  else if (tb.G <= 16) launch_rorr_p<16>(tb, ch, worklist, wcount, wcap, exact_only, st);
  else launch_rorr_p<32>(tb, ch, worklist, wcount, wcap, exact_only, st);
}

void rt_launch_expand_opacity(const DevTables &tb, const DevChunk &ch, int icol, double *tau,
                              double *w0, double *g, double *tband, cudaStream_t st) {
  expand_opacity_kernel<<<592, 256, 0, st>>>(tb, ch, icol, tau, w0, g, tband);
}

void rt_launch_tau_band(const DevTables &tb, const DevChunk &ch, int is_ir, double *out,
                        cudaStream_t st) {
  const int s = is_ir ? tb.ir_start : tb.sol_start, e = is_ir ? tb.ir_end : tb.sol_end;
  tau_band_kernel<<<592, 256, 0, st>>>(tb, ch, s, e - s + 1, out);
}
