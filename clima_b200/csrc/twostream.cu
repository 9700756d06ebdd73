// twostream.cu -- Toon-89 two-stream solver (IR + solar), Planck source, g/zenith weighting.
//
// Replaces two_stream_ir / two_stream_solar / tridiag
// (src/radtran/clima_radtran_twostream.f90:10-316) and the per-bin loop of `radiate`
// (src/radtran/clima_radtran_radiate.f90:50-158) for a batch of columns.
//
// Mapping: one THREAD per (column, channel bin ll, g-point): the 2nz x 2nz tridiagonal system is
// solved by the same Thomas recurrence, in the same row order, as `tridiag` (:297-316), so the
// result tracks the reference to rounding.  The P (= ngauss rounded to a power of two) threads of
// one (column, bin) are adjacent lanes: tau/w0 loads are full 8*G-byte lines and the g-weighted
// sum (radiate.f90:122-126) is a sub-warp shuffle reduction.
//
// The forward sweep streams layer by layer and keeps only the previous layer in registers; what
// the back substitution needs per layer (c', d' of its two rows, Gamma, exp(-lambda tau), C+(tau),
// C-(tau)) goes to a global scratch laid out [layer][item][resident thread] (coalesced 8-byte
// accesses; the kernel is persistent so the scratch is sized by resident threads, not by work).
//
// Work skipped because it is provably invariant:
//   * duplicated layers (same slot as the previous layer) reuse gamma/lambda/Gamma/exp;
//   * solar: the matrix (A,B,D) does not depend on the zenith angle (A.5) and elimination, back
//     substitution and flux reconstruction are all linear in the right-hand side, so the zenith
//     quadrature is taken on the RHS terms themselves (sum_z w_z C+-_z, sum_z w_z direct_z) and ONE
//     Thomas solve runs on the summed system; per zenith angle only exp(-tau/u0), C+- and four
//     multiply-adds per layer remain;
//   * the direct beam exp(-tauc/u0) is carried as a running product (one exp per layer and zenith
//     instead of two, :75-76).
#include "rt_device.cuh"
#include "rt_kernels.h"

namespace {

constexpr int TS_THREADS = 128;
#ifndef TS_IR_BLOCKS
#define TS_IR_BLOCKS 5
#endif
#ifndef TS_SOL_BLOCKS
#define TS_SOL_BLOCKS 3
#endif
// forward-sweep unroll (register renaming instead of the moves that rotate the layer state).
// Measured r1r per 1024 columns: IR 4.08 (1) / 4.17 (2) ms; solar 5.64 (1) / 5.42 (2) / 5.67 (3) ms.  Re-measured r2f
// on the r1s kernels, per 4096 columns: IR 14.46 (1) / 13.63 (2) / 13.69 (3) / 13.66 (4) ms; solar 20.74 (1) / 19.57 (2) /
// 20.01 (3) / 20.70 (4) ms
#ifndef TS_FWD_UNROLL_IR
#define TS_FWD_UNROLL_IR 2
#endif
#ifndef TS_FWD_UNROLL_SOL
#define TS_FWD_UNROLL_SOL 2
#endif
constexpr int TS_FWD_UNROLL_IR_N = TS_FWD_UNROLL_IR, TS_FWD_UNROLL_SOL_N = TS_FWD_UNROLL_SOL;
// Scratch per layer, [layer][item][resident thread] followed by [layer][resident group][slot]:
//   IR    4 numbers per thread (alpha, beta, Bu, Bd) + per (column, bin) group the g-weighted sums of
//         the terms that do not multiply the solution (sum w Au, sum w Ad): the kernel is HBM-bound,
//         the two extra shuffle reductions of the forward sweep buy 31 % less traffic
//         (r1r per 1024 columns: 4.01 -> 3.64 ms);
//   solar 7 numbers per thread (alpha, beta, Au, Bu, Ad, Bd, K).  Group sums in this kernel's forward
//         sweep cost more than the traffic they save (r1r per 1024 columns, zenith-folded sweep:
//         4.88 ms as is, 5.30 with sum w K per group, 5.26 with Au/Ad/K per group, 5.15 at 4 blocks/SM).
// cache hints on the per-thread scratch (0 none, 1 streaming stores, 2 streaming stores + loads, 3 loads only).
// Measured r2f per 4096 columns (IR / solar ms): 0: 14.55 / 19.73, 1: 14.58 / 19.78, 2: 15.10 / 21.19, 3: 15.07 / 21.17
#ifndef TS_SCR_HINT
#define TS_SCR_HINT 0
#endif
__device__ __forceinline__ void scr_st(double *p, double v) {
#if TS_SCR_HINT == 1 || TS_SCR_HINT == 2
  __stcs(p, v);
#else
  *p = v;
#endif
}
__device__ __forceinline__ double scr_ld(const double *p) {
#if TS_SCR_HINT >= 2
  return __ldcs(p);
#else
  return *p;
#endif
}
constexpr int IR_TITEMS = 4, IR_GITEMS = 2;
#ifndef TS_SOL_PRESUM
#define TS_SOL_PRESUM 0  // 0: all 7 per thread; 1: sum w K per group; 2: sum w Au, sum w Ad, sum w K per group
#endif
constexpr int SOL_TITEMS = TS_SOL_PRESUM == 2 ? 4 : (TS_SOL_PRESUM == 1 ? 6 : 7);
constexpr int SOL_GITEMS = TS_SOL_PRESUM == 2 ? 4 : 1;  // group slots: (Au, Ad, K, -) or (K)

// forward sweep: tau/w0 lines this many layers ahead into L2 (0 = off).  Measured r1q per 1024 columns:
// IR 5.16 (0) / 5.21 (6) ms; solar 6.94 (0) / 6.75 (6) / 6.93 (12) ms
// Re-measured r2g (IR forward sweep unrolled by 2), per 4096 columns: IR 13.71 (0) / 13.19 (2) / 13.19 (4) / 13.37 (6-8) /
// 13.65 (12) ms; solar 21.42 (0) / 19.6 (6) / 20.82 (12) ms
#ifndef TS_FWD_L2_PREFETCH_IR
#define TS_FWD_L2_PREFETCH_IR 4
#endif
#ifndef TS_FWD_L2_PREFETCH_SOL
#define TS_FWD_L2_PREFETCH_SOL 6
#endif
// layers ahead of the back substitution (0 = off).  Measured r1r per 1024 columns (IR / solar):
// 0: 4.30 / 5.51, 2: 4.08 / 5.39, 4: 4.01 / 5.38, 6: 4.04 / 5.40, 8: 4.08 / 5.42, 16: 4.87 / 5.98 ms
#ifndef TS_L2_PREFETCH_IR
#define TS_L2_PREFETCH_IR 4
#endif
#ifndef TS_L2_PREFETCH_SOL
#define TS_L2_PREFETCH_SOL 4
#endif
// The scratch of a persistent two-stream grid (~0.9 GB) is far larger than L2: by the time the back
// substitution reads a layer it has left the cache.  The register prefetch reaches two layers
// (~100 cycles of arithmetic); an L2 prefetch a few layers further ahead costs no registers.
__device__ __forceinline__ void l2_prefetch(const double *p) {
  asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
}

template <int P>
__device__ __forceinline__ double group_sum(double v) {
#pragma unroll
  for (int m = P / 2; m >= 1; m >>= 1) v += __shfl_xor_sync(0xffffffffu, v, m);
  return v;
}

// Two sums over the P lanes of a group for the price of one and a third: the first exchange halves
// the data (even lanes keep a and receive the partner's a, odd lanes keep b), the remaining steps
// reduce one value.  Returns sum(a) in the even lanes of the group and sum(b) in the odd lanes.
template <int P>
__device__ __forceinline__ double group_sum2(const double a, const double b, const int gl) {
  const bool odd = (gl & 1) != 0;
  double v = (odd ? b : a) + __shfl_xor_sync(0xffffffffu, odd ? a : b, 1);
#pragma unroll
  for (int m = P / 2; m >= 2; m >>= 1) v += __shfl_xor_sync(0xffffffffu, v, m);
  return v;
}

__global__ void planck_kernel(const DevTables tb, const DevChunk ch, double *bpl) {
  const int nz = ch.nz, nw_ir = tb.ir_end - tb.ir_start + 1;
  const size_t total = (size_t)ch.ncol * nw_ir * (nz + 1);
  for (size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x; t < total;
       t += (size_t)gridDim.x * blockDim.x) {
    const int n = (int)(t % (nz + 1));
    const int ll = (int)((t / (nz + 1)) % nw_ir);
    const int c = (int)(t / ((size_t)(nz + 1) * nw_ir));
    const int l = tb.ir_start + ll;
    const double avg_freq = 0.5 * (tb.freq[l] + tb.freq[l + 1]);  // radiate.f90:64
    const double T = (n == nz) ? ch.T_surface[c] : ch.T[(size_t)c * nz + (nz - 1 - n)];  // :65-69
    bpl[t] = rt_planck(avg_freq, T);
  }
}

// ------------------------------------------------------------------------------------------ IR
template <int P>
__global__ void __launch_bounds__(TS_THREADS, TS_IR_BLOCKS)
twostream_ir_kernel(const __grid_constant__ DevTables tb, const __grid_constant__ DevParams pr,
                    const __grid_constant__ DevChunk ch, const __grid_constant__ DevFlux fx,
                    const int ocol_bcast) {
  constexpr int GPB = TS_THREADS / P;  // (col,bin) groups per block
  const int G = tb.G, nw = tb.nw, nz = ch.nz, cap = ch.nslot_cap;
  const int nw_ir = tb.ir_end - tb.ir_start + 1;
  const long long ngroups = (long long)ch.ncol * nw_ir;
  const size_t nres = (size_t)gridDim.x * TS_THREADS;
  const size_t rtid = (size_t)blockIdx.x * TS_THREADS + threadIdx.x;
  double *scr = fx.scratch + rtid;
  const int gl = threadIdx.x % P;
  const size_t ngres = nres / P;
  // this lane's slot of its group's sums (even lanes: the upward term, odd lanes: the downward one)
  double *gscr = fx.scratch + (size_t)IR_TITEMS * nz * nres + (rtid / P) * IR_GITEMS + (gl & 1);
  const int gq = gl < G ? gl : G - 1;
  const double wg = gl < G ? tb.wbin[gq] : 0.0;
  const double norm = 2.0 * RT_PI * 0.5;  // twostream.f90:185-186 (u1 = 0.5)
  const bool hard = pr.has_hard_surface != 0;
  const double tmin = pr.ir_tau_min;

  for (long long w = blockIdx.x; w * GPB < ngroups; w += gridDim.x) {
    long long gid = w * GPB + threadIdx.x / P;
    const bool gactive = gid < ngroups;
    if (!gactive) gid = ngroups - 1;
    const int c = (int)(gid / nw_ir), ll = (int)(gid % nw_ir);
    const int l = tb.ir_start + ll;
    const int oc = ocol_bcast ? 0 : c;
    const size_t ob = ((size_t)oc * nw + l) * cap;
    const double *tau_p = ch.tau + ob * G + gq;
    const double *w0_p = ch.w0 + ob * G + gq;
    const double *gt_p = ch.gt + ob;
    const int *sol = ch.slot_of_layer + (size_t)oc * nz;
    const double *bp = fx.bpl + ((size_t)c * nw_ir + ll) * (nz + 1);
    const double emis = pr.emiss[ll];
    const double Rsfc = hard ? 1.0 - emis : 0.0;

    // ---- forward sweep (top layer first)
    double e1, e2, e3, e4, cpb, cmb, cgam, wrk;  // layer i
    double tau_i, g12;                           // layer i
    double cprev, dprev;                         // c', d' of row 2i
    double cp0_top;
    int slot = sol[nz - 1];
    {
      const double tau = tau_p[(size_t)slot * G], w0 = w0_p[(size_t)slot * G], gt = gt_p[slot];
      const double gam1 = 2.0 - w0 * (1.0 + gt), gam2 = w0 * (1.0 - gt);  // :195-196
      const double lam = sqrt(gam1 * gam1 - gam2 * gam2);                 // :199
      cgam = gam2 * rt_rcp_fast(gam1 + lam);                              // :200
      wrk = rt_exp_neg(-lam * tau);                                       // :206
      e1 = 1.0 + cgam * wrk; e2 = 1.0 - cgam * wrk; e3 = cgam + wrk; e4 = cgam - wrk;
      g12 = rt_rcp_fast(gam1 + gam2);
      tau_i = tau;
      double b0, b1;  // :216-227
      const double Bt = bp[0], Bb = bp[1];
      if (tau <= tmin) { b0 = 0.5 * (Bt + Bb); b1 = 0.0; } else { b0 = Bt; b1 = (Bb - b0) * rt_rcp_fast(tau); }
      const double cp0 = norm * (b0 + b1 * g12);
      cpb = norm * (b0 + b1 * (tau + g12));
      const double cm0 = norm * (b0 + b1 * (-g12));
      cmb = norm * (b0 + b1 * (tau - g12));
      cp0_top = cp0;
      // row 1 (:251-254) and its Thomas step (:302-303)
      const double re1 = rt_rcp_fast(e1);
      cprev = (-e2) * re1;
      dprev = (0.0 - cm0) * re1;
    }
    // Raw inputs are prefetched one layer ahead (slot index two layers ahead, so that the dependent
    // tau/w0/g loads can be issued a full iteration before they are needed): the sweep was
    // latency-bound on the chain slot -> tau -> coefficients (r1e profile: 53 % long-scoreboard).
    auto lslot = [&](int i) { return sol[nz - 1 - (i < nz ? i : nz - 1)]; };  // slot of layer i (clamped)
    double Ta = 0.0, Tb = 0.0;
    int s1 = lslot(1), s2 = lslot(2);
    double t1 = tau_p[(size_t)s1 * G], w1 = w0_p[(size_t)s1 * G], g1 = gt_p[s1];
    double B1 = bp[nz >= 1 ? 1 : 0], B2 = bp[nz >= 2 ? 2 : nz];
#pragma unroll TS_FWD_UNROLL_IR_N
    for (int i = 0; i < nz; i++) {
      // issue the loads of layer i+2 / level i+3
      const int s3 = lslot(i + 3);
#if TS_FWD_L2_PREFETCH_IR > 0
      {  // tau / w0 lines of a layer further down, into L2
        const int sp = lslot(i + 3 + TS_FWD_L2_PREFETCH_IR);
        l2_prefetch(tau_p + (size_t)sp * G);
        l2_prefetch(w0_p + (size_t)sp * G);
      }
#endif
      const double t2 = tau_p[(size_t)s2 * G], w2 = w0_p[(size_t)s2 * G], g2 = gt_p[s2];
      const double B3 = bp[i + 3 <= nz ? i + 3 : nz];
      double c1, d1;              // row 2i+1 (0-based)
      double cn = 0.0, dn = 0.0;  // row 2i+2
      double e1n = 0, e2n = 0, e3n = 0, e4n = 0, cpbn = 0, cmbn = 0, cgamn = 0, wrkn = 0, taun = 0, g12n = 0;
      if (i < nz - 1) {
        const int slotn = s1;
        if (slotn == slot) {
          cgamn = cgam; wrkn = wrk; e1n = e1; e2n = e2; e3n = e3; e4n = e4; taun = tau_i; g12n = g12;
        } else {
          const double tau = t1, w0 = w1, gt = g1;
          const double gam1 = 2.0 - w0 * (1.0 + gt), gam2 = w0 * (1.0 - gt);
          const double lam = sqrt(gam1 * gam1 - gam2 * gam2);
          cgamn = gam2 * rt_rcp_fast(gam1 + lam);
          wrkn = rt_exp_neg(-lam * tau);
          e1n = 1.0 + cgamn * wrkn; e2n = 1.0 - cgamn * wrkn; e3n = cgamn + wrkn; e4n = cgamn - wrkn;
          g12n = rt_rcp_fast(gam1 + gam2);
          taun = tau;
          slot = slotn;
        }
        double b0, b1;
        const double Bt = B1, Bb = B2;
        if (taun <= tmin) { b0 = 0.5 * (Bt + Bb); b1 = 0.0; } else { b0 = Bt; b1 = (Bb - b0) * rt_rcp_fast(taun); }
        const double cp0n = norm * (b0 + b1 * g12n);
        cpbn = norm * (b0 + b1 * (taun + g12n));
        const double cm0n = norm * (b0 + b1 * (-g12n));
        cmbn = norm * (b0 + b1 * (taun - g12n));
        // even Fortran row l=2i (:266-270) -> 0-based row 2i+1
        {
          const double A = e2n * e1 - e3 * e4n;
          const double B = e2 * e2n - e4 * e4n;
          const double D = e1n * e4n - e2n * e3n;
          const double E = e2n * (cp0n - cpb) - e4n * (cm0n - cmb);
          const double den = B - A * cprev;
          const double rden = rt_rcp_fast(den);  // one reciprocal per row
          c1 = D * rden;
          d1 = (E - A * dprev) * rden;
        }
        // odd Fortran row l=2i+1 (:257-262) -> 0-based row 2i+2
        {
          const double A = e2 * e3 - e4 * e1;
          const double B = e1 * e1n - e3 * e3n;
          const double D = e3 * e4n - e1 * e2n;
          const double E = e3 * (cp0n - cpb) + e1 * (cmb - cm0n);
          const double den = B - A * c1;
          const double rden = rt_rcp_fast(den);
          cn = D * rden;
          dn = (E - A * d1) * rden;
        }
      } else {
        // last row (:272-275); surface terms (:236-247)
        double Ssfc;
        if (hard) {
          Ssfc = emis * RT_PI * bp[nz];
        } else {
          double b1_bot = 0.0;
          if (!(tau_i <= tmin)) b1_bot = (bp[nz] - bp[nz - 1]) * rt_rcp_fast(tau_i);
          Ssfc = RT_PI * (bp[nz] + 0.5 * b1_bot);
        }
        const double A = e1 - Rsfc * e3;
        const double B = e2 - Rsfc * e4;
        const double E = Ssfc - cpb + Rsfc * cmb;
        c1 = 0.0;
        d1 = (E - A * dprev) * rt_rcp_fast(B - A * cprev);
      }
      // What the back substitution needs from layer i, folded into 6 numbers.  With xn = x of row
      // 2i+2 (the solution entering the layer from below), Thomas (:311-313) gives
      //   y2 = x1 = d1 - c1*xn ,  y1 = x0 = d0 - c0*x1 = alpha + beta*xn
      // and the fluxes (:288-293) are affine in xn as well:
      //   fup = y1*e1 + y2*e2 + cpb = Au + Bu*xn ,  fdn = y1*e3 + y2*e4 + cmb = Ad + Bd*xn.
      {
        const double c0 = cprev, d0 = dprev;
        const double Qu = e2 - c0 * e1, Qd = e4 - c0 * e3;
        const double al = d0 - c0 * d1, be = c0 * c1;
        double *s = scr + (size_t)i * IR_TITEMS * nres;
        scr_st(s, al); scr_st(s + nres, be);
        scr_st(s + 2 * nres, -(Qu * c1)); scr_st(s + 3 * nres, -(Qd * c1));
        // Au, Ad enter the fluxes only through their g-weighted sums (radiate.f90:122-126)
        const double SA = group_sum2<P>(((d0 * e1 + cpb) + Qu * d1) * wg, ((d0 * e3 + cmb) + Qd * d1) * wg, gl);
        if (gl < 2) gscr[(size_t)i * IR_GITEMS * ngres] = SA;
        if (i == 0) {  // top-of-atmosphere upward flux (:288): y1*e3 - y2*e4 + cp0 = Ta + Tb*xn
          Ta = (al * e3 - d1 * e4) + cp0_top;
          Tb = be * e3 + c1 * e4;
        }
      }
      e1 = e1n; e2 = e2n; e3 = e3n; e4 = e4n; cpb = cpbn; cmb = cmbn; cgam = cgamn; wrk = wrkn;
      tau_i = taun; g12 = g12n; cprev = cn; dprev = dn;
      s1 = s2; s2 = s3; t1 = t2; w1 = w2; g1 = g2; B1 = B2; B2 = B3;
    }

    // ---- back substitution (:311-313) + fluxes (:288-293) + g-weights (radiate.f90:122-126)
    double *fup_o = fx.fup_a_ir + ((size_t)c * nw_ir + ll) * (nz + 1);
    double *fdn_o = fx.fdn_a_ir + ((size_t)c * nw_ir + ll) * (nz + 1);
    const bool writer = gactive && gl == 0;
    double *const out_o = (gl & 1) ? fdn_o : fup_o;  // lane 0 writes fup, lane 1 fdn
    const bool writer2 = gactive && gl < 2;
    double xnext = 0.0, xtop = 0.0;
    // the scratch values of a layer do not depend on the recurrence: keep the loads two layers
    // ahead of the arithmetic (rotating register buffers, loop unrolled by 3)
    double q[3][IR_TITEMS + 1];
    auto ld = [&](double (&b)[IR_TITEMS + 1], int i) {
      const int ic = i > 0 ? i : 0;
      const double *s = scr + (size_t)ic * IR_TITEMS * nres;
#pragma unroll
      for (int t = 0; t < IR_TITEMS; t++) b[t] = scr_ld(s + (size_t)t * nres);
      b[IR_TITEMS] = gscr[(size_t)ic * IR_GITEMS * ngres];
      if (TS_L2_PREFETCH_IR > 0 && i >= TS_L2_PREFETCH_IR) {
        const double *pf = scr + (size_t)(i - TS_L2_PREFETCH_IR) * IR_TITEMS * nres;
#pragma unroll
        for (int t = 0; t < IR_TITEMS; t++) l2_prefetch(pf + (size_t)t * nres);
        l2_prefetch(gscr + (size_t)(i - TS_L2_PREFETCH_IR) * IR_GITEMS * ngres);
      }
    };
    auto bk = [&](const double (&b)[IR_TITEMS + 1], int i) {
      // fup = Au + Bu*xn, fdn = Ad + Bd*xn summed over g: even lanes get fup, odd lanes fdn
      const double wx = wg * xnext;
      const double R = group_sum2<P>(b[2] * wx, b[3] * wx, gl) + b[IR_TITEMS];
      if (writer2) out_o[nz - 1 - i] = R;
      xtop = xnext;
      xnext = b[0] + b[1] * xnext;
    };
    ld(q[0], nz - 1);
    ld(q[1], nz - 2);
    for (int i = nz - 1; i >= 0; i -= 3) {
      ld(q[2], i - 2);
      bk(q[0], i);
      if (i - 1 >= 0) { ld(q[0], i - 3); bk(q[1], i - 1); }
      if (i - 2 >= 0) { ld(q[1], i - 4); bk(q[2], i - 2); }
    }
    {
      const double fup = Ta + Tb * xtop;  // :288
      const double su = group_sum<P>(fup * wg);
      if (writer) { fup_o[nz] = su; fdn_o[nz] = 0.0; }
    }
  }
}

// --------------------------------------------------------------------------------------- solar
template <int P, int ZC>
__global__ void __launch_bounds__(TS_THREADS, TS_SOL_BLOCKS)
twostream_solar_kernel(const __grid_constant__ DevTables tb, const __grid_constant__ DevParams pr,
                       const __grid_constant__ DevChunk ch, const __grid_constant__ DevFlux fx,
                       const int ocol_bcast, const int zbeg, const int accumulate) {
  constexpr int GPB = TS_THREADS / P;
  const int G = tb.G, nw = tb.nw, nz = ch.nz, cap = ch.nslot_cap;
  const int nw_sol = tb.sol_end - tb.sol_start + 1;
  const long long ngroups = (long long)ch.ncol * nw_sol;
  const size_t nres = (size_t)gridDim.x * TS_THREADS;
  const size_t rtid = (size_t)blockIdx.x * TS_THREADS + threadIdx.x;
  double *scr = fx.scratch + rtid;
  const int gl = threadIdx.x % P;
  const size_t ngres = nres / P;
  double *gscr = fx.scratch + (size_t)SOL_TITEMS * nz * nres + (rtid / P) * SOL_GITEMS;  // the group's sum w K
  const int gq = gl < G ? gl : G - 1;
  const double wg = gl < G ? tb.wbin[gq] : 0.0;
  const double sq3 = sqrt(3.0);
  const double u1i = 1.0 / (1.0 / sqrt(3.0));  // (1/u1), u1 = 1/sqrt(3)  (:30,:135)

  double u0[ZC], iu0[ZC], iu02[ZC], wz[ZC];
#pragma unroll
  for (int z = 0; z < ZC; z++) {
    const bool on = zbeg + z < pr.nzen;
    u0[z] = on ? pr.zen_u[zbeg + z] : 1.0;
    wz[z] = on ? pr.zen_w[zbeg + z] : 0.0;
    iu0[z] = 1.0 / u0[z];
    iu02[z] = 1.0 / (u0[z] * u0[z]);
  }

  for (long long w = blockIdx.x; w * GPB < ngroups; w += gridDim.x) {
    long long gid = w * GPB + threadIdx.x / P;
    const bool gactive = gid < ngroups;
    if (!gactive) gid = ngroups - 1;
    const int c = (int)(gid / nw_sol), ll = (int)(gid % nw_sol);
    const int l = tb.sol_start + ll;
    const int oc = ocol_bcast ? 0 : c;
    const size_t ob = ((size_t)oc * nw + l) * cap;
    const double *tau_p = ch.tau + ob * G + gq;
    const double *w0_p = ch.w0 + ob * G + gq;
    const double *gt_p = ch.gt + ob;
    const int *sol = ch.slot_of_layer + (size_t)oc * nz;
    const double Rsfc = pr.albedo[ll];

    // layer i state (zenith independent)
    double e1, e2, e3, e4, cgam, wrk;
    // per-zenith layer coefficients (reused by duplicated layers) and running state
    double fp[ZC], fm[ZC], ex[ZC];  // facp/denom, facm/denom, exp(-tau'/u0)
    double aw[ZC];                  // w_z * exp(-tauc/u0) at the bottom of layer i (= top of layer i+1)
    double CPB, CMB;                // sum_z w_z C+-(bottom of layer i)
    double cprev;
    double CP0_top = 0.0, FDN_top = 0.0, AM_top = 0.0;
    int slot = -1;

    auto layer_coeffs = [&](const double tau_in, const double w0_in, const double gt_in) {
      // delta-Eddington scaling (:38-40)
      const double f = 1.0 - w0_in * gt_in * gt_in;
      const double tau = tau_in * f;
      const double w0 = w0_in * (1.0 - gt_in * gt_in) * rt_rcp_fast(f);
      const double gt = gt_in * rt_rcp_fast(1.0 + gt_in);
      // quadrature coefficients (:43-46)
      const double gam1 = sq3 * (2.0 - w0 * (1.0 + gt)) / 2.0;
      const double gam2 = sq3 * w0 * (1.0 - gt) / 2.0;
      const double lam = sqrt(gam1 * gam1 - gam2 * gam2);  // :50
      cgam = gam2 * rt_rcp_fast(gam1 + lam);               // :51
      wrk = rt_exp_neg(-lam * tau);                        // :56
      e1 = 1.0 + cgam * wrk; e2 = 1.0 - cgam * wrk; e3 = cgam + wrk; e4 = cgam - wrk;
#pragma unroll
      for (int z = 0; z < ZC; z++) {
        const double gam3 = (1.0 - sq3 * gt * u0[z]) / 2.0;  // :45
        const double gam4 = 1.0 - gam3;
        const double facp = w0 * ((gam1 - iu0[z]) * gam3 + gam4 * gam2);  // :76
        const double facm = w0 * ((gam1 + iu0[z]) * gam4 + gam2 * gam3);  // :77
        const double rden = rt_rcp_fast(lam * lam - iu02[z]);             // :80
        fp[z] = facp * rden;
        fm[z] = facm * rden;
        ex[z] = rt_exp_neg(-tau * iu0[z]);
      }
    };

    // ---- top layer and row 1
    slot = sol[nz - 1];
    layer_coeffs(tau_p[(size_t)slot * G], w0_p[(size_t)slot * G], gt_p[slot]);
    // raw inputs prefetched one layer ahead, slot index two layers ahead (see the IR kernel)
    auto lslot = [&](int i) { return sol[nz - 1 - (i < nz ? i : nz - 1)]; };
    int s1 = lslot(1), s2 = lslot(2);
    double t1 = tau_p[(size_t)s1 * G], w1 = w0_p[(size_t)s1 * G], g1 = gt_p[s1];
    // The whole solve is linear in the right-hand side and the matrix does not depend on the zenith
    // angle: the quadrature sum_z w_z is taken on the RHS terms (C+-, direct beam) BEFORE the
    // elimination, so one Thomas recurrence runs on D' = sum_z w_z d'_z.
    const double re1_top = rt_rcp_fast(e1);
    double Dprev;  // sum_z w_z d'_z of row 2i
    {
      double CM0 = 0.0;
      CPB = 0.0; CMB = 0.0;
#pragma unroll
      for (int z = 0; z < ZC; z++) {
        CP0_top += wz[z] * fp[z];  // exp(-tauc(1)/u0) = 1 at the top
        CM0 += wz[z] * fm[z];
        aw[z] = wz[z] * ex[z];
        CPB += aw[z] * fp[z];
        CMB += aw[z] * fm[z];
        FDN_top += wz[z] * u0[z];  // direct(1) = u0*Fs_pi (:73)
        AM_top += wz[z];           // direct(1)/u0
      }
      Dprev = (0.0 - CM0) * re1_top;  // rows :93-96, Thomas :302-303
    }
    cprev = (-e2) * re1_top;
    double Ta = 0.0, Tb = 0.0;

#pragma unroll TS_FWD_UNROLL_SOL_N
    for (int i = 0; i < nz; i++) {
      const int s3 = lslot(i + 3);
#if TS_FWD_L2_PREFETCH_SOL > 0
      {  // tau / w0 lines of a layer further down, into L2
        const int sp = lslot(i + 3 + TS_FWD_L2_PREFETCH_SOL);
        l2_prefetch(tau_p + (size_t)sp * G);
        l2_prefetch(w0_p + (size_t)sp * G);
      }
#endif
      const double t2 = tau_p[(size_t)s2 * G], w2 = w0_p[(size_t)s2 * G], g2 = gt_p[s2];
      // layer i ("o"ld) values; layer_coeffs() overwrites e1..e4/cgam/wrk/fp/fm/ex with layer i+1
      const double e1o = e1, e2o = e2, e3o = e3, e4o = e4;
      double c1, cn = 0.0, D1, Dn = 0.0;
      double CPBn = 0.0, CMBn = 0.0, DIR = 0.0, DIRU = 0.0;
#pragma unroll
      for (int z = 0; z < ZC; z++) {
        DIR += aw[z] * u0[z];  // direct(i+1) = u0*etb(i) (:84)
        DIRU += aw[z];         // direct(i+1)/u0
      }
      if (i < nz - 1) {
        const int slotn = s1;
        if (slotn != slot) {
          layer_coeffs(t1, w1, g1);
          slot = slotn;
        }
        // even Fortran row l=2i (:107-112) -> 0-based row 2i+1 ; (e1..e4 = layer i+1)
        const double A1 = e2 * e1o - e3o * e4, B1 = e2o * e2 - e4o * e4, Dd1 = e1 * e4 - e2 * e3;
        const double r1 = rt_rcp_fast(B1 - A1 * cprev);
        c1 = Dd1 * r1;
        // odd Fortran row l=2i+1 (:99-104) -> 0-based row 2i+2
        const double A2 = e2o * e3o - e4o * e1o, B2 = e1o * e1 - e3o * e3, Dd2 = e3o * e4 - e1o * e2;
        const double r2 = rt_rcp_fast(B2 - A2 * c1);
        cn = Dd2 * r2;
        // zenith sums of C+-(top of layer i+1) and, after the layer's attenuation, C+-(its bottom)
        double CP0n = 0.0, CM0n = 0.0;
#pragma unroll
        for (int z = 0; z < ZC; z++) {
          CP0n += aw[z] * fp[z];
          CM0n += aw[z] * fm[z];
          aw[z] *= ex[z];
          CPBn += aw[z] * fp[z];
          CMBn += aw[z] * fm[z];
        }
        const double E1 = e2 * (CP0n - CPB) - e4 * (CM0n - CMB);
        D1 = (E1 - A1 * Dprev) * r1;
        const double E2 = e3o * (CP0n - CPB) + e1o * (CMB - CM0n);
        Dn = (E2 - A2 * D1) * r2;
      } else {
        const double A = e1o - Rsfc * e3o, B = e2o - Rsfc * e4o;  // :114-117
        const double r1 = rt_rcp_fast(B - A * cprev);
        c1 = 0.0;
        const double E = Rsfc * DIR - CPB + Rsfc * CMB;  // Ssfc = Rsfc*u0*etb (:90), summed over z
        D1 = (E - A * Dprev) * r1;
      }
      // folded back-substitution data of layer i (see the IR kernel): with xn entering from below,
      //   y1 = alpha + beta*xn, fup = Au + Bu*xn, fdn (incl. direct beam) = Ad + Bd*xn,
      //   amean = (1/u1)*(fup + fdn) + K,  K = direct/u0 - (1/u1)*direct   (:135-148)
      {
        const double c0 = cprev, d0 = Dprev;
        const double Qu = e2o - c0 * e1o, Qd = e4o - c0 * e3o;
        const double al = d0 - c0 * D1, be = c0 * c1;
        double *s = scr + (size_t)i * SOL_TITEMS * nres;
        scr_st(s, al); scr_st(s + nres, be);
        const double Au = (d0 * e1o + CPB) + Qu * D1, Ad = ((d0 * e3o + CMB) + Qd * D1) + DIR;
        const double Kk = DIRU - u1i * DIR;
        // Au, Ad and K enter the outputs only through their g-weighted sums (radiate.f90:122-126)
#if TS_SOL_PRESUM == 2
        s[2 * nres] = -(Qu * c1); s[3 * nres] = -(Qd * c1);
        const double SA = group_sum2<P>(Au * wg, Ad * wg, gl);
        const double SK = group_sum<P>(Kk * wg);
        if (gl < 3) gscr[(size_t)i * SOL_GITEMS * ngres + gl] = gl == 2 ? SK : SA;
#else
        scr_st(s + 2 * nres, Au); scr_st(s + 3 * nres, -(Qu * c1));
        scr_st(s + 4 * nres, Ad); scr_st(s + 5 * nres, -(Qd * c1));
#if TS_SOL_PRESUM == 1
        const double SK = group_sum<P>(Kk * wg);
        if (gl == 0) gscr[(size_t)i * SOL_GITEMS * ngres] = SK;
#else
        scr_st(s + 6 * nres, Kk);
#endif
#endif
        if (i == 0) {  // :143  y1*e3 - y2*e4 + cp0 = Ta + Tb*xn
          Ta = (al * e3o - D1 * e4o) + CP0_top;
          Tb = be * e3o + c1 * e4o;
        }
      }
      cprev = cn;
      Dprev = Dn;
      CPB = CPBn; CMB = CMBn;
      s1 = s2; s2 = s3; t1 = t2; w1 = w2; g1 = g2;
    }

    // ---- ONE back substitution for the zenith-weighted solution, fluxes (:135-148), g-weights
    const size_t obase = ((size_t)c * nw_sol + ll) * (nz + 1);
    double *fup_o = fx.fup_a_sol + obase, *fdn_o = fx.fdn_a_sol + obase, *am_o = fx.amean_sol + obase;
    // scale factors applied after the g/zenith sums (radiate.f90:167-180)
    const double ps = pr.photons_sol[ll] * pr.photon_scale_factor;  // clima_radtran.f90:302
    const double avg_freq = 0.5 * (tb.freq[l] + tb.freq[l + 1]);
    const double avg_wavl = 1.0e9 * RT_C_LIGHT / avg_freq;
    const double am1 = (avg_freq / avg_wavl);
    const double am2 = (avg_wavl / (RT_PLANK * RT_C_LIGHT * 1.0e16));
    const double dwl = (tb.wavl[l + 1] - tb.wavl[l]);
    // R: sum over g of the upward (even lanes) / downward (odd lanes) flux; sa: of the mean intensity
    double *const out_o = (gl & 1) ? fdn_o : fup_o;  // lane 0 writes fup and amean, lane 1 fdn
    auto emit = [&](const int ig, const double R, const double sa) {
      if (gactive && gl < 2) {
        double v = R * ps * pr.diurnal_fac;
        if (accumulate) v += out_o[ig];
        out_o[ig] = v;
        if (gl == 0) {
          double m = sa * ps * pr.diurnal_fac;
          m = m * am1;
          m = m * am2 * dwl;
          if (accumulate) m += am_o[ig];
          am_o[ig] = m;
        }
      }
    };
    double xnext = 0.0, xtop = 0.0;
    // scratch loads two layers ahead of the arithmetic (rotating register buffers, unrolled by 3)
    constexpr int NB = SOL_TITEMS + 2;  // + the group's sums (this lane's parity of Au/Ad; K)
    double q[3][NB];
    auto ld = [&](double (&b)[NB], int i) {
      const int ic = i > 0 ? i : 0;
      const double *s = scr + (size_t)ic * SOL_TITEMS * nres;
#pragma unroll
      for (int t = 0; t < SOL_TITEMS; t++) b[t] = scr_ld(s + (size_t)t * nres);
      const double *gs = gscr + (size_t)ic * SOL_GITEMS * ngres;
      b[SOL_TITEMS] = TS_SOL_PRESUM == 2 ? gs[gl & 1] : 0.0;
      b[SOL_TITEMS + 1] = TS_SOL_PRESUM == 2 ? gs[2] : (TS_SOL_PRESUM == 1 ? gs[0] : 0.0);
      if (TS_L2_PREFETCH_SOL > 0 && i >= TS_L2_PREFETCH_SOL) {
        const double *pf = scr + (size_t)(i - TS_L2_PREFETCH_SOL) * SOL_TITEMS * nres;
#pragma unroll
        for (int t = 0; t < SOL_TITEMS; t++) l2_prefetch(pf + (size_t)t * nres);
        if (TS_SOL_PRESUM) l2_prefetch(gscr + (size_t)(i - TS_L2_PREFETCH_SOL) * SOL_GITEMS * ngres);
      }
    };
    auto bk = [&](const double (&b)[NB], int i) {
      // fup = Au + Bu*xn (:146), fdn = Ad + Bd*xn (:147), amean = (1/u1)(fup + fdn) + K (:137-139),
      // all three summed over g
#if TS_SOL_PRESUM == 2
      const double wx = wg * xnext;
      const double R = group_sum2<P>(b[2] * wx, b[3] * wx, gl) + b[SOL_TITEMS];
#else
      const double R = group_sum2<P>((b[2] + b[3] * xnext) * wg, (b[4] + b[5] * xnext) * wg, gl);
#endif
      const double Ro = __shfl_xor_sync(0xffffffffu, R, 1);
      const double SK = TS_SOL_PRESUM ? b[SOL_TITEMS + 1] : group_sum<P>(b[SOL_TITEMS - 1] * wg);
      emit(nz - 1 - i, R, u1i * (R + Ro) + SK);
      xtop = xnext;
      xnext = b[0] + b[1] * xnext;
    };
    ld(q[0], nz - 1);
    ld(q[1], nz - 2);
    for (int i = nz - 1; i >= 0; i -= 3) {
      ld(q[2], i - 2);
      bk(q[0], i);
      if (i - 1 >= 0) { ld(q[0], i - 3); bk(q[1], i - 1); }
      if (i - 2 >= 0) { ld(q[1], i - 4); bk(q[2], i - 2); }
    }
    {
      const double up = Ta + Tb * xtop;  // :143
      emit(nz, group_sum2<P>(up * wg, FDN_top * wg, gl), group_sum<P>((u1i * up + AM_top) * wg));  // :135, :144
    }
  }
}

// ------------------------------------------------------------------------- frequency integration
// radiate.f90:184-192 + clima_radtran.f90:286-289,316 + TOA_fluxes (:339-340).
// One thread per (column, level); the sum over bins runs in the reference's order.
__global__ void integrate_kernel(const DevTables tb, const DevChunk ch, const DevFlux fx,
                                 const int compute_solar, const int sol_bcast) {
  const int nz = ch.nz, L = nz + 1;
  const int nw_ir = tb.ir_end - tb.ir_start + 1, nw_sol = tb.sol_end - tb.sol_start + 1;
  const size_t total = (size_t)ch.ncol * L;
  for (size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x; t < total;
       t += (size_t)gridDim.x * blockDim.x) {
    const int i = (int)(t % L);
    const int c = (int)(t / L);
    double up = 0.0, dn = 0.0;
    for (int ll = 0; ll < nw_ir; ll++) {
      const int l = tb.ir_start + ll;
      const double dfreq = tb.freq[l] - tb.freq[l + 1];
      up = up + fx.fup_a_ir[((size_t)c * nw_ir + ll) * L + i] * dfreq;
      dn = dn + fx.fdn_a_ir[((size_t)c * nw_ir + ll) * L + i] * dfreq;
    }
    fx.fup_n_ir[t] = up;
    fx.fdn_n_ir[t] = dn;
    double sup, sdn;
    if (compute_solar) {
      sup = 0.0; sdn = 0.0;
      for (int ll = 0; ll < nw_sol; ll++) {
        const int l = tb.sol_start + ll;
        const double dfreq = tb.freq[l] - tb.freq[l + 1];
        sup = sup + fx.fup_a_sol[((size_t)c * nw_sol + ll) * L + i] * dfreq;
        sdn = sdn + fx.fdn_a_sol[((size_t)c * nw_sol + ll) * L + i] * dfreq;
      }
      fx.fup_n_sol[t] = sup;
      fx.fdn_n_sol[t] = sdn;
    } else {
      // stale solar fluxes of the previous call (clima_radtran.f90:286-289)
      const size_t ts = sol_bcast ? (size_t)i : t;
      sup = fx.fup_n_sol[ts];
      sdn = fx.fdn_n_sol[ts];
      if (sol_bcast && c > 0) { fx.fup_n_sol[t] = sup; fx.fdn_n_sol[t] = sdn; }
    }
    fx.f_total[t] = (sdn - sup) + (dn - up);
    if (i == nz) {
      fx.ISR[c] = (sdn - sup);
      fx.OLR[c] = -(dn - up);
      if (!(isfinite(sdn) && isfinite(sup) && isfinite(dn) && isfinite(up))) atomicOr(fx.status + c, 2);
    }
  }
}

}  // namespace

// Resident blocks per SM of the persistent two-stream kernels (grid = SMs x this, so that every
// block of the grid is resident at once and the grid-stride loop has no half-empty second wave).
template <int P>
static void occ_p(int nzen, int *ir, int *sol) {
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(ir, twostream_ir_kernel<P>, TS_THREADS, 0);
  int a = 0, b = 0, c = 0;
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&a, twostream_solar_kernel<P, 4>, TS_THREADS, 0);
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, twostream_solar_kernel<P, 2>, TS_THREADS, 0);
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&c, twostream_solar_kernel<P, 1>, TS_THREADS, 0);
  (void)nzen;
  *sol = a < b ? (a < c ? a : c) : (b < c ? b : c);
}
void rt_twostream_occupancy(int G, int nzen, int *ir_blocks_per_sm, int *sol_blocks_per_sm) {
  if (G <= 8) occ_p<8>(nzen, ir_blocks_per_sm, sol_blocks_per_sm);
  else if (G <= 16) occ_p<16>(nzen, ir_blocks_per_sm, sol_blocks_per_sm);
  else occ_p<32>(nzen, ir_blocks_per_sm, sol_blocks_per_sm);
  if (*ir_blocks_per_sm < 1) *ir_blocks_per_sm = 1;
  if (*sol_blocks_per_sm < 1) *sol_blocks_per_sm = 1;
}

int rt_twostream_block_threads() { return TS_THREADS; }
// doubles per layer and resident thread the scratch must hold (per-thread items + the group slots, P >= 8)
int rt_twostream_items_per_layer() { return 8; }

void rt_launch_planck(const DevTables &tb, const DevChunk &ch, const DevFlux &fx, cudaStream_t st) {
  planck_kernel<<<1184, 256, 0, st>>>(tb, ch, fx.bpl);
}

void rt_launch_twostream_ir(const DevTables &tb, const DevParams &pr, const DevChunk &ch,
                            const DevFlux &fx, int ocol_bcast, int nblocks, cudaStream_t st) {
  if (tb.G <= 8) twostream_ir_kernel<8><<<nblocks, TS_THREADS, 0, st>>>(tb, pr, ch, fx, ocol_bcast);
  else if (tb.G <= 16) twostream_ir_kernel<16><<<nblocks, TS_THREADS, 0, st>>>(tb, pr, ch, fx, ocol_bcast);
  else twostream_ir_kernel<32><<<nblocks, TS_THREADS, 0, st>>>(tb, pr, ch, fx, ocol_bcast);
}

template <int P>
static void launch_solar_p(const DevTables &tb, const DevParams &pr, const DevChunk &ch,
                           const DevFlux &fx, int ocol_bcast, int nblocks, cudaStream_t st) {
  // zenith angles in passes of at most 4 (register budget); later passes accumulate
  int zbeg = 0;
  while (zbeg < pr.nzen) {
    const int left = pr.nzen - zbeg;
    const int acc = zbeg > 0;
    if (left >= 4) {
      twostream_solar_kernel<P, 4><<<nblocks, TS_THREADS, 0, st>>>(tb, pr, ch, fx, ocol_bcast, zbeg, acc);
      zbeg += 4;
    } else if (left >= 2) {
      twostream_solar_kernel<P, 2><<<nblocks, TS_THREADS, 0, st>>>(tb, pr, ch, fx, ocol_bcast, zbeg, acc);
      zbeg += 2;
    } else {
      twostream_solar_kernel<P, 1><<<nblocks, TS_THREADS, 0, st>>>(tb, pr, ch, fx, ocol_bcast, zbeg, acc);
      zbeg += 1;
    }
  }
}

void rt_launch_twostream_solar(const DevTables &tb, const DevParams &pr, const DevChunk &ch,
                               const DevFlux &fx, int ocol_bcast, int nblocks, cudaStream_t st) {
  if (tb.G <= 8) launch_solar_p<8>(tb, pr, ch, fx, ocol_bcast, nblocks, st);
  else if (tb.G <= 16) launch_solar_p<16>(tb, pr, ch, fx, ocol_bcast, nblocks, st);
  else launch_solar_p<32>(tb, pr, ch, fx, ocol_bcast, nblocks, st);
}

void rt_launch_integrate(const DevTables &tb, const DevParams &pr, const DevChunk &ch,
                         const DevFlux &fx, int compute_solar, int sol_bcast, cudaStream_t st) {
  (void)pr;
  const size_t total = (size_t)ch.ncol * (ch.nz + 1);
  int nb = (int)((total + 255) / 256);
  if (nb > 4736) nb = 4736;
  integrate_kernel<<<nb, 256, 0, st>>>(tb, ch, fx, compute_solar, sol_bcast);
}
