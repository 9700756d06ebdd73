// jacobian_delta.cu -- N1: the frozen-opacity finite-difference Jacobian as a DELTA solve with a
// factorisation shared by all columns.
//
// Replaces, for the columns of clima_rt_jacobian_batch, the loop of AdiabatClimate_jacobian_from_base
// (src/adiabat/clima_adiabat_solve.f90:798-820): every perturbed column is `radiate` with
// compute_opacity=.false., compute_solar=.false., i.e. two_stream_ir (clima_radtran_twostream.f90:156-295)
// on the SAME (tau, w0, g) with another Planck vector, then f_total(2i-1) (:731-733).
//
// All columns share the 2nz x 2nz tridiagonal matrix (:251-275); the right-hand side is linear in the
// Planck values, and a finite-difference column differs from the base in 2-4 levels.  With the
// top-down factors c'_r and the bottom-up factors a''_r of the same matrix,
//     x_r = kappa_r (d'_r - c'_r d''_(r+1)),     kappa_r = 1 / (1 - c'_r a''_(r+1)),
// where d' is the top-down and d'' the bottom-up elimination of the right-hand side.  For a right-hand
// side that is non-zero only in rows r0..r1:  d'_r = 0 above the window and the chain
// d'_r = phi_r d'_(r-1) below it; d''_r = 0 below and d''_r = psi_r d''_(r+1) above.  The solution
// everywhere therefore follows from a window solve of <= 16 rows and two chains of one multiply per
// row -- no per-row storage per column, no exp, no sqrt (experiments/jacobian_delta_proto.py).
//
//   jac_factor_kernel   once per base column: one block per (IR bin, g); coefficients e1..e4 in
//                       parallel over the layers, the two factorisations as two serial sweeps;
//                       writes psi, c'kappa, phi, kappa, A, 1/den, D, 1/den'' per row and e1..e4,
//                       1/(gam1+gam2), tau, 1/tau (thin-layer flag), e3-e1, e4-e2 per layer
//   jac_window_kernel   per column: first / last radiative level whose temperature differs from the base
//   jac_delta_kernel    lane = column, a warp works through a list of (bin, g) units: every factor load is
//                       the same address for the whole warp (one transaction); accumulates
//                       sum_bins dnu sum_g w_g (dfdn - dfup) at the nz/2 output levels in shared memory
//   jac_reduce_kernel   adds the per-split partial sums in a fixed order and the base column's f_total
// The result differs from re-solving every column only by rounding (the difference of two solves is
// formed analytically instead of numerically).
#include <algorithm>

#include "rt_device.cuh"
#include "rt_kernels.h"

namespace {

constexpr int JR = 8;      // per-row arrays
constexpr int JL = 9;      // per-layer arrays
constexpr int MAXLEV = 5;  // changed levels a column may span (2 for a model layer, 4 for the top one, 1 for the surface)
constexpr int MAXW = 14;   // rows of the window: 2 (lev1 - lev0) + 6 <= 14 for MAXLEV levels
enum { R_PSI = 0, R_CK, R_PHI, R_KAP, R_A, R_RDEN, R_D, R_RDEN2 };
enum { L_E1 = 0, L_E2, L_E3, L_E4, L_G12, L_TAU, L_RTAU, L_U, L_V };

constexpr int JG = 4;      // per-group arrays (a group = the 4 rows of an even layer and the odd layer below it)
enum { G_GU = 0, G_PU, G_GD, G_PD };
__host__ __device__ inline size_t jac_stride(int nz) { return (size_t)JR * 2 * nz + (size_t)JL * nz + (size_t)JG * (nz / 2); }

__global__ void __launch_bounds__(64)
jac_factor_kernel(const __grid_constant__ DevTables tb, const __grid_constant__ DevParams pr,
                  const __grid_constant__ DevChunk ch, double *__restrict__ factor) {
  extern __shared__ double sm[];  // e1..e4 [4][nz], then A, B, D [3][2nz]
  const int nz = ch.nz, n = 2 * nz, G = tb.G, nw = tb.nw, cap = ch.nslot_cap;
  const int ll = blockIdx.x / G, g = blockIdx.x % G;
  const int l = tb.ir_start + ll;
  double *fr = factor + (size_t)blockIdx.x * jac_stride(nz);  // rows [JR][n]
  double *fl = fr + (size_t)JR * n;                            // layers [JL][nz]
  double *e1 = sm, *e2 = sm + nz, *e3 = sm + 2 * nz, *e4 = sm + 3 * nz;
  double *A = sm + 4 * nz, *B = A + n, *D = B + n;
  const size_t ob = ((size_t)0 * nw + l) * cap;  // opacity of column 0 (the base)
  const int *sol = ch.slot_of_layer;
  const bool hard = pr.has_hard_surface != 0;
  const double Rsfc = hard ? 1.0 - pr.emiss[ll] : 0.0;
  // ---- coefficients, parallel over layers (top-first layer i = ground-first layer nz-1-i)
  for (int i = threadIdx.x; i < nz; i += blockDim.x) {
    const int slot = sol[nz - 1 - i];
    const double tau = ch.tau[(ob + slot) * G + g], w0 = ch.w0[(ob + slot) * G + g], gt = ch.gt[ob + slot];
    const double gam1 = 2.0 - w0 * (1.0 + gt), gam2 = w0 * (1.0 - gt);  // :195-196
    const double lam = sqrt(gam1 * gam1 - gam2 * gam2);                 // :199
    const double cgam = gam2 / (gam1 + lam);                            // :200
    const double wrk = exp(-lam * tau);                                 // :206
    const double a1 = 1.0 + cgam * wrk, a2 = 1.0 - cgam * wrk, a3 = cgam + wrk, a4 = cgam - wrk;
    e1[i] = a1; e2[i] = a2; e3[i] = a3; e4[i] = a4;
    fl[L_E1 * nz + i] = a1; fl[L_E2 * nz + i] = a2; fl[L_E3 * nz + i] = a3; fl[L_E4 * nz + i] = a4;
    fl[L_G12 * nz + i] = 1.0 / (gam1 + gam2);
    fl[L_TAU * nz + i] = tau;
    fl[L_RTAU * nz + i] = (tau <= pr.ir_tau_min) ? -1.0 : 1.0 / tau;  // < 0: thin-layer guard (:216-227)
    fl[L_U * nz + i] = a3 - a1;
    fl[L_V * nz + i] = a4 - a2;
  }
  __syncthreads();
  // ---- the matrix (:251-275), parallel over interfaces
  for (int i = threadIdx.x; i < nz - 1; i += blockDim.x) {
    A[2 * i + 1] = e2[i + 1] * e1[i] - e3[i] * e4[i + 1];
    B[2 * i + 1] = e2[i] * e2[i + 1] - e4[i] * e4[i + 1];
    D[2 * i + 1] = e1[i + 1] * e4[i + 1] - e2[i + 1] * e3[i + 1];
    A[2 * i + 2] = e2[i] * e3[i] - e4[i] * e1[i];
    B[2 * i + 2] = e1[i] * e1[i + 1] - e3[i] * e3[i + 1];
    D[2 * i + 2] = e3[i] * e4[i + 1] - e1[i] * e2[i + 1];
  }
  if (threadIdx.x == 0) {
    A[0] = 0.0; B[0] = e1[0]; D[0] = -e2[0];
    A[n - 1] = e1[nz - 1] - Rsfc * e3[nz - 1]; B[n - 1] = e2[nz - 1] - Rsfc * e4[nz - 1]; D[n - 1] = 0.0;
  }
  __syncthreads();
  // ---- two factorisations, one serial sweep each (two warps)
  if (threadIdx.x == 0) {  // top-down: c'_r, 1/den_r, phi_r = -A_r/den_r   (tridiag, :297-316)
    double cp = 0.0;
    for (int r = 0; r < n; r++) {
      const double rden = 1.0 / (B[r] - A[r] * cp);
      cp = D[r] * rden;
      fr[R_RDEN * n + r] = rden;
      fr[R_PHI * n + r] = -A[r] * rden;
      fr[R_A * n + r] = A[r];
      fr[R_D * n + r] = D[r];
      B[r] = cp;  // reuse: c'_r for the kappa pass
    }
  }
  if (threadIdx.x == 32) {  // bottom-up: a''_r, 1/den''_r, psi_r = -D_r/den''_r
    double ap = 0.0;
    for (int r = n - 1; r >= 0; r--) {
      const double b = (r == 0) ? e1[0] : ((r == n - 1) ? (e2[nz - 1] - Rsfc * e4[nz - 1])
                       : ((r & 1) ? (e2[(r - 1) / 2] * e2[(r + 1) / 2] - e4[(r - 1) / 2] * e4[(r + 1) / 2])
                                  : (e1[r / 2 - 1] * e1[r / 2] - e3[r / 2 - 1] * e3[r / 2])));
      const double rden2 = 1.0 / (b - D[r] * ap);
      ap = A[r] * rden2;
      fr[R_RDEN2 * n + r] = rden2;
      fr[R_PSI * n + r] = -D[r] * rden2;
      fr[R_CK * n + r] = ap;  // a''_r for now
    }
  }
  __syncthreads();
  // ---- kappa_r = 1/(1 - c'_r a''_(r+1)),  c'kappa stored negated: x_r = kappa d' + ck d''_(r+1)
  __threadfence_block();
  for (int r = threadIdx.x; r < n; r += blockDim.x) {
    const double cp = B[r];
    const double ap1 = (r + 1 < n) ? fr[R_CK * n + r + 1] : 0.0;
    const double kap = 1.0 / (1.0 - cp * ap1);
    fr[R_KAP * n + r] = kap;
    A[r] = -cp * kap;  // staged: R_CK still holds a'' that neighbours read
  }
  __syncthreads();
  for (int r = threadIdx.x; r < n; r += blockDim.x) fr[R_CK * n + r] = A[r];
  __syncthreads();
  // ---- groups.  Outputs are wanted at the bottom of the ODD layers t = 2m+1 only; the two chains pass
  // through the 4 rows 4m..4m+3 (even layer 2m, odd layer 2m+1) with ONE multiply and add ONE product:
  //   upward (rows 4m+3 .. 4m, d'' entering from below):  dnet += GU[m] d'',  d'' *= PU[m]
  //   downward (rows 4m .. 4m+3, d' entering from above): dnet += GD[m] d',   d'  *= PD[m]
  double *fg = fl + (size_t)JL * nz;
  for (int m = threadIdx.x; m < nz / 2; m += blockDim.x) {
    const int t = 2 * m + 1, r = 4 * m;
    const double U = fl[L_U * nz + t], V = fl[L_V * nz + t];
    const double p0 = fr[R_PSI * n + r], p1 = fr[R_PSI * n + r + 1], p2 = fr[R_PSI * n + r + 2], p3 = fr[R_PSI * n + r + 3];
    fg[G_GU * (nz / 2) + m] = A[r + 3] * V + A[r + 2] * p3 * U;  // A[] holds -c'kappa
    fg[G_PU * (nz / 2) + m] = ((p3 * p2) * p1) * p0;
    const double f0 = fr[R_PHI * n + r], f1 = fr[R_PHI * n + r + 1], f2 = fr[R_PHI * n + r + 2], f3 = fr[R_PHI * n + r + 3];
    const double k2 = fr[R_KAP * n + r + 2], k3 = fr[R_KAP * n + r + 3];
    fg[G_GD * (nz / 2) + m] = ((f0 * f1) * f2) * (k2 * U + f3 * k3 * V);
    fg[G_PD * (nz / 2) + m] = ((f0 * f1) * f2) * f3;
  }
}

// first / last top-first radiative LEVEL whose temperature differs from the base column
// (level t < nz: temperature of top-first layer t; level nz: T_surface; radiate.f90:65-69), from the
// model-grid profiles (copy_atm_to_radiative_grid, clima_adiabat.f90:729-773: layer i -> 2i, 2i+1, two
// ghost copies of the top layer).  win[2c] > win[2c+1]: unchanged column.  flag: a window too wide.
__global__ void jac_window_kernel(const double *__restrict__ Tm, const double *__restrict__ Ts,
                                  const double *__restrict__ baseT, int nzm, int ncol, int *__restrict__ win,
                                  int *__restrict__ flag) {
  const int nzr = 2 * nzm + 2;
  for (int c = blockIdx.x * blockDim.x + threadIdx.x; c < ncol; c += gridDim.x * blockDim.x) {
    int lo = nzr + 1, hi = -1;
    for (int j = 0; j < nzr; j++) {  // ground-first radiative layer j
      const int i = j < 2 * nzm ? j / 2 : nzm - 1;
      if (Tm[(size_t)c * nzm + i] != baseT[j]) {
        const int t = nzr - 1 - j;
        lo = min(lo, t); hi = max(hi, t);
      }
    }
    if (Ts[c] != baseT[nzr]) { lo = min(lo, nzr); hi = max(hi, nzr); }
    win[2 * c] = lo; win[2 * c + 1] = hi;
    if (hi >= lo && hi - lo + 1 > MAXLEV) atomicOr(flag, 1);
  }
}

// lane = column; the warp walks units u = u0, u0 + ustep, ... (u = ll*G + g).  partial:
// [split][col][nzm+1] sums of dnu * w_g * (dfdn - dfup) at model level i (ground-first).
#ifndef JAC_MIN_BLOCKS
#define JAC_MIN_BLOCKS 3  // measured r2d, 1024 stacked Jacobians: 2 -> 26.2 ms, 3 -> 20.4, 4 -> 22.1, 6 -> 25.7
#endif
__global__ void __launch_bounds__(128, JAC_MIN_BLOCKS)
jac_delta_kernel(const __grid_constant__ DevTables tb, const __grid_constant__ DevParams pr, const int nzm,
                 const int ncol, const double *__restrict__ Tm, const double *__restrict__ Ts,
                 const double *__restrict__ baseT, const int *__restrict__ win,
                 const double *__restrict__ factor, double *__restrict__ partial, const int nsplit) {
  extern __shared__ double acc[];  // [warp][nzm+1][32] sums, then [warp][2][4*(nzm+1)] staged group factors
  const int nz = 2 * nzm + 2, n = 2 * nz, G = tb.G;
  const int nw_ir = tb.ir_end - tb.ir_start + 1, nunits = nw_ir * G;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, wpb = blockDim.x >> 5;
  const int cgrp = blockIdx.x, split = blockIdx.y;
  const int c = cgrp * 32 + lane;
  const bool live = c < ncol;
  const int cc = live ? c : ncol - 1;
  double *my = acc + (size_t)warp * (nzm + 1) * 32 + lane;
  // the group factors of a unit (GU, PU, GD, PD: 4 (nzm+1) doubles) are staged through shared memory by one
  // coalesced warp load, two units ahead in two buffers: the chains were bound on the latency of ~200
  // dependent-issue uniform global loads per unit (ncu r2d: long scoreboard 6.3 warps per issue, 25 % issue)
  const int ng4 = 4 * (nzm + 1);
  double *sgw = acc + (size_t)wpb * (nzm + 1) * 32 + (size_t)warp * 2 * ng4;
  for (int i = 0; i <= nzm; i++) my[i * 32] = 0.0;
  const int lev0 = win[2 * cc], lev1 = win[2 * cc + 1];
  const bool changed = lev1 >= lev0 && (lev1 - lev0 + 1) <= MAXLEV;
  const bool hard = pr.has_hard_surface != 0;
  // temperatures of the changed levels, perturbed and base
  double Tp[MAXLEV], Tb[MAXLEV];
#pragma unroll
  for (int k = 0; k < MAXLEV; k++) {
    const int t = lev0 + k;
    double a = 1.0, b = 1.0;
    if (changed && t <= lev1) {
      if (t == nz) { a = Ts[cc]; b = baseT[nz]; }
      else {
        const int j = nz - 1 - t;
        const int i = j < 2 * nzm ? j / 2 : nzm - 1;
        a = Tm[(size_t)cc * nzm + i]; b = baseT[j];
      }
    }
    Tp[k] = a; Tb[k] = b;
  }
  // affected layers la0..la1 (Bt or Bb changed), window rows r0..r1
  // an unchanged column gets an empty window (r0 > r1): every delta below is zero
  const int la0 = changed ? max(0, lev0 - 1) : 1, la1 = changed ? min(nz - 1, lev1) : 0;
  const int r0 = changed ? max(0, 2 * (lev0 - 2) + 1) : n, r1 = changed ? min(n - 1, 2 * lev1 + 2) : -1;

  // this warp's contiguous range of units (g fastest: the Planck deltas are per bin, reused over g)
  const int nwarps = nsplit * wpb, wid = split * wpb + warp;
  const int ubeg = (int)((long long)nunits * wid / nwarps), uend = (int)((long long)nunits * (wid + 1) / nwarps);
  const int ngrp = nz / 2;
  int ll_cur = -1;
  double dB[MAXLEV + 2];  // dB[k] = delta B(level lev0 - 1 + k): zero pads on both sides
#pragma unroll
  for (int k = 0; k < MAXLEV + 2; k++) dB[k] = 0.0;
  auto stage = [&](int u, int buf) {  // group factors of unit u -> buffer buf (coalesced)
    const double *src = factor + (size_t)u * jac_stride(nz) + (size_t)JR * n + (size_t)JL * nz;
    double *dst = sgw + buf * ng4;
    for (int k = lane; k < ng4; k += 32) dst[k] = src[k];
  };
  if (ubeg < uend) stage(ubeg, 0);
  __syncwarp();
  for (int u = ubeg; u < uend; u++) {
    const int ll = u / G, g = u - ll * G;
    const int l = tb.ir_start + ll;
    const double *fr = factor + (size_t)u * jac_stride(nz);
    const double *fl = fr + (size_t)JR * n;
    const double *fg = sgw + ((u - ubeg) & 1) * ng4;
    if (u + 1 < uend) stage(u + 1, (u + 1 - ubeg) & 1);  // in flight while this unit's window is solved
    const double wgt = (tb.freq[l] - tb.freq[l + 1]) * tb.wbin[g];  // integrate_kernel x radiate.f90:122-126
    if (ll != ll_cur) {
      ll_cur = ll;
      const double avg_freq = 0.5 * (tb.freq[l] + tb.freq[l + 1]);  // radiate.f90:64
#pragma unroll
      for (int k = 0; k < MAXLEV; k++)
        dB[k + 1] = (changed && lev0 + k <= lev1) ? rt_planck(avg_freq, Tp[k]) - rt_planck(avg_freq, Tb[k]) : 0.0;
    }
    // delta source terms of the affected layers (linear in the Planck values, :216-233)
    double dcp0[MAXLEV + 1], dcpb[MAXLEV + 1], dcm0[MAXLEV + 1], dcmb[MAXLEV + 1];
#pragma unroll
    for (int k = 0; k <= MAXLEV; k++) {
      const int i = lev0 - 1 + k;  // layer: Bt = level i, Bb = level i + 1
      double p0 = 0.0, pb = 0.0, m0 = 0.0, mb = 0.0;
      if (i >= la0 && i <= la1) {
        const double dBt = dB[k], dBb = dB[k + 1];
        const double rtau = fl[L_RTAU * nz + i], tau = fl[L_TAU * nz + i], g12 = fl[L_G12 * nz + i];
        const double b0 = rtau < 0.0 ? 0.5 * (dBt + dBb) : dBt;
        const double b1 = rtau < 0.0 ? 0.0 : (dBb - dBt) * rtau;
        p0 = RT_PI * (b0 + b1 * g12); pb = RT_PI * (b0 + b1 * (tau + g12));
        m0 = RT_PI * (b0 - b1 * g12); mb = RT_PI * (b0 + b1 * (tau - g12));
        if (i & 1) my[((nz - 1 - i) >> 1) * 32] += wgt * (mb - pb);  // source term of an output layer (:288-293)
      }
      dcp0[k] = p0; dcpb[k] = pb; dcm0[k] = m0; dcmb[k] = mb;
    }
    auto lay = [&](const double (&arr)[MAXLEV + 1], int i) -> double {  // delta of layer i (0 outside)
      const int k = i - (lev0 - 1);
      double v = 0.0;
#pragma unroll
      for (int q = 0; q <= MAXLEV; q++) v = (q == k) ? arr[q] : v;
      return v;
    };
    // Output: model level i (ground-first) = bottom of top-first layer t = nz - 1 - 2i (odd t):
    //   d(fdn - fup) = x_2t (e3 - e1) + x_(2t+1) (e4 - e2) + (dcmb - dcpb)      (:288-293)
    // every ROW of an odd layer contributes x_r * (r even ? U : V).
    auto addrow = [&](int r, double x) {
      const int t = r >> 1;
      if (t & 1) my[((nz - 1 - t) >> 1) * 32] += wgt * x * fl[((r & 1) ? L_V : L_U) * nz + t];
    };
    // ---- window: delta right-hand side, top-down d', bottom-up d'', the window's own rows
    double dE[MAXW], dp[MAXW];
    double prev = 0.0;
#pragma unroll
    for (int w = 0; w < MAXW; w++) {
      const int r = r0 + w;
      double e = 0.0;
      if (r <= r1) {
        if (r == 0) {
          e = -lay(dcm0, 0);
        } else if (r == n - 1) {
          double dS;
          const double dBs = (lev1 == nz) ? dB[nz - lev0 + 1] : 0.0;  // level nz = the surface
          if (hard) dS = pr.emiss[ll] * RT_PI * dBs;
          else {
            const double rtau = fl[L_RTAU * nz + nz - 1];
            const double dBm = (lev0 <= nz - 1 && nz - 1 <= lev1) ? dB[nz - 1 - lev0 + 1] : 0.0;
            dS = RT_PI * (dBs + 0.5 * (rtau < 0.0 ? 0.0 : (dBs - dBm) * rtau));
          }
          const double Rs = hard ? 1.0 - pr.emiss[ll] : 0.0;
          e = dS - lay(dcpb, nz - 1) + Rs * lay(dcmb, nz - 1);
        } else {
          const int i = (r - 1) >> 1;  // interface between layers i and i + 1
          const double dp0 = lay(dcp0, i + 1) - lay(dcpb, i), dm0 = lay(dcm0, i + 1) - lay(dcmb, i);
          if (r & 1) e = fl[L_E2 * nz + i + 1] * dp0 - fl[L_E4 * nz + i + 1] * dm0;
          else e = fl[L_E3 * nz + i] * dp0 - fl[L_E1 * nz + i] * dm0;
        }
        prev = (e - fr[R_A * n + r] * prev) * fr[R_RDEN * n + r];
      }
      dE[w] = e;
      dp[w] = prev;  // keeps d'_(r1) beyond the window
    }
    double d2 = 0.0;  // d''_(r+1) while walking the window upward; ends as d''_(r0)
#pragma unroll
    for (int w = MAXW - 1; w >= 0; w--) {
      const int r = r0 + w;
      if (r <= r1) {
        addrow(r, fr[R_KAP * n + r] * dp[w] + fr[R_CK * n + r] * d2);
        d2 = (dE[w] - fr[R_D * n + r] * d2) * fr[R_RDEN2 * n + r];
      }
    }
    double d1 = prev;  // d'_(r1)
    // ---- above the window: single rows up to the next group boundary, then whole groups
    {
      int r = r0 - 1;
      for (; r >= 0 && (r & 3) != 3; r--) {
        addrow(r, fr[R_CK * n + r] * d2);
        d2 *= fr[R_PSI * n + r];
      }
      const int mtop = changed ? (r >> 2) : -1;  // first whole group (r = 4m+3), -1: none
      const int mmax = __reduce_max_sync(0xffffffffu, mtop);
#pragma unroll 4
      for (int m = mmax; m >= 0; m--) {
        const double gu = fg[G_GU * ngrp + m], pu = fg[G_PU * ngrp + m];
        if (m <= mtop) {
          my[(nzm - m) * 32] += wgt * gu * d2;
          d2 *= pu;
        }
      }
    }
    // ---- below the window
    {
      int r = r1 + 1;
      for (; r < n && (r & 3) != 0; r++) {
        d1 *= fr[R_PHI * n + r];
        addrow(r, fr[R_KAP * n + r] * d1);
      }
      const int mbot = changed ? (r >> 2) : ngrp;  // first whole group (r = 4m), ngrp: none
      const int mmin = __reduce_min_sync(0xffffffffu, mbot);
#pragma unroll 4
      for (int m = mmin; m < ngrp; m++) {
        const double gd = fg[G_GD * ngrp + m], pd = fg[G_PD * ngrp + m];
        if (m >= mbot) {
          my[(nzm - m) * 32] += wgt * gd * d1;
          d1 *= pd;
        }
      }
    }
    __syncwarp();  // every lane is done with this unit's buffer; the next unit's is complete
  }
  __syncwarp();
  // cross-warp sum in warp order, then out
  __syncthreads();
  if (warp == 0 && live) {
    for (int i = 0; i <= nzm; i++) {
      double s = 0.0;
      for (int w = 0; w < wpb; w++) s += acc[((size_t)w * (nzm + 1) + i) * 32 + lane];
      partial[((size_t)split * ncol + c) * (nzm + 1) + i] = s;
    }
  }
}

}  // namespace

size_t rt_jac_factor_doubles(const DevTables &tb, int nz) {
  return (size_t)(tb.ir_end - tb.ir_start + 1) * tb.G * jac_stride(nz);
}

// factor: rt_jac_factor_doubles(tb, ch.nz) doubles; ch holds the base column's opacity in column 0
void rt_launch_jac_factor(const DevTables &tb, const DevParams &pr, const DevChunk &ch, double *factor,
                          cudaStream_t st) {
  const int nunits = (tb.ir_end - tb.ir_start + 1) * tb.G;
  const size_t smem = (size_t)10 * ch.nz * sizeof(double);
  static bool attr_set[64] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  if (smem > 48 * 1024 && (dev < 0 || dev >= 64 || !attr_set[dev])) {
    if (cudaFuncSetAttribute(jac_factor_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) == cudaSuccess &&
        dev >= 0 && dev < 64)
      attr_set[dev] = true;
  }
  jac_factor_kernel<<<nunits, 64, smem, st>>>(tb, pr, ch, factor);
}

int rt_jac_splits(const DevTables &tb, int ncol, int num_sms) {
  // enough warps to fill the GPU: column groups x splits x 4 warps >= ~8 warps per SM
  const int groups = (ncol + 31) / 32;
  const int nunits = (tb.ir_end - tb.ir_start + 1) * tb.G;
  int s = (num_sms * 8 + groups * 4 - 1) / (groups * 4);
  s = s < 1 ? 1 : s;
  const int smax = (nunits + 3) / 4;
  return s > smax ? smax : s;
}

// Tm (nzm, ncol) / Ts (ncol): perturbed model-grid temperatures; baseT: nzr radiative-grid layer
// temperatures (ground-first) + T_surface; win: 2*ncol ints; partial: nsplit*ncol*(nzm+1) doubles;
// flag: set when a column's changed levels span more than the window holds (the caller falls back)
void rt_launch_jac_delta(const DevTables &tb, const DevParams &pr, int nzm, int ncol, const double *Tm,
                         const double *Ts, const double *baseT, const double *factor, int *win, int *flag,
                         double *partial, int nsplit, cudaStream_t st) {
  cudaMemsetAsync(flag, 0, sizeof(int), st);
  jac_window_kernel<<<(ncol + 127) / 128, 128, 0, st>>>(Tm, Ts, baseT, nzm, ncol, win, flag);
  const size_t smem = (size_t)4 * ((nzm + 1) * 32 + 2 * 4 * (nzm + 1)) * sizeof(double);
  static bool attr_set[64] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  if (smem > 48 * 1024 && (dev < 0 || dev >= 64 || !attr_set[dev])) {
    if (cudaFuncSetAttribute(jac_delta_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) == cudaSuccess &&
        dev >= 0 && dev < 64)
      attr_set[dev] = true;
  }
  dim3 grid((ncol + 31) / 32, nsplit);
  jac_delta_kernel<<<grid, 128, smem, st>>>(tb, pr, nzm, ncol, Tm, Ts, baseT, win, factor, partial, nsplit);
}

namespace {
// f(i, c) = base f_total(2i) + sum over splits of the partial sums, in split order
__global__ void jac_reduce_kernel(const double *__restrict__ partial, const double *__restrict__ basef,
                                  double *__restrict__ out, int nzm, int ncol, int nsplit) {
  const size_t total = (size_t)ncol * (nzm + 1);
  for (size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x; t < total; t += (size_t)gridDim.x * blockDim.x) {
    const int i = (int)(t % (nzm + 1));
    double s = 0.0;
    for (int k = 0; k < nsplit; k++) s += partial[(size_t)k * total + t];
    out[t] = basef[2 * i] + s;
  }
}
}  // namespace

void rt_launch_jac_reduce(const double *partial, const double *basef, double *out, int nzm, int ncol, int nsplit,
                          cudaStream_t st) {
  const size_t total = (size_t)ncol * (nzm + 1);
  jac_reduce_kernel<<<(unsigned)std::min<size_t>((total + 255) / 256, 1184), 256, 0, st>>>(partial, basef, out, nzm, ncol, nsplit);
}
