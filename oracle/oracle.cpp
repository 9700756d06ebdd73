// oracle.cpp -- CPU restatement of Clima's radiative-transfer hot path.
//
// *** TEST INFRASTRUCTURE ONLY. ***  This file is the parity oracle and the timed CPU baseline.
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
// load it.  Nothing under clima_b200/ links, imports or calls it; the product path is CUDA only.
//
// PARITY STATUS: "parity unpinned".  The reference (Fortran 2008) cannot be compiled in this
// image (no Fortran compiler, no fypp/HDF5, CPM dependencies need network) and its own tests
// hold no numeric assertions or golden vectors for this path (tests/test_radtran.f90:67-81 only
// prints).  Two routines live in the un-vendored dependency Nicholaswogan/futils v0.1.14
// (src/dependencies/CMakeLists.txt:35-42): `mrgrnk` (ORDERPACK merge-sort ranking) and `rebin`;
// they are restated here from their published semantics (stable ascending rank; overlap-weighted
// bin average).  The oracle is pinned instead by (1) an independent NumPy/SciPy restatement
// (oracle/oracle_np.py), (2) closed-form limits and conservation properties (tests/), and
// (3) frozen golden vectors of its own output (tests/golden/).
//
// Every function cites the reference file:line it follows.  Loop order and operation order follow
// the Fortran statement by statement; build with -ffp-contract=off for the parity build (gfortran
// on x86-64 without -march does not contract), -O3 -march=native -fopenmp for the timed baseline.

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <string>
#include <vector>

#ifdef _OPENMP
#include <omp.h>
#endif

#include "../include/clima_rt.h"

namespace {

// src/clima_const.f90:9-21
constexpr double k_boltz_si = 1.380649e-23;
constexpr double plank = 6.62607004e-34;
constexpr double c_light = 299792458.0;
constexpr double pi = 3.14159265358979323846e0;

// src/radtran/clima_radtran_types.f90:9-11
constexpr double max_w0 = 0.99999;
constexpr double max_gt = 0.999999;
constexpr double tau_min = 1.0e-20;

// ---------------------------------------------------------------------------------------------
// src/dependencies/linear_interpolation_module.F90:364-501 (dintrv_real).  The cached `ilo` only
// accelerates the search; the returned bracket is fully specified by :348-350:
//   x <  xt(1)          -> (1,2)
//   xt(i) <= x < xt(i+1) -> (i,i+1)
//   xt(n) <= x          -> (n-1,n)
// Returned index is 0-based ileft; iright = ileft+1.
inline int dintrv(const double *xt, int n, double x) {
  if (x >= xt[n - 1]) return n - 2;
  if (x < xt[0]) return 0;
  int lo = 0, hi = n - 1;  // xt[lo] <= x < xt[hi]
  while (hi - lo > 1) {
    int mid = (lo + hi) / 2;
    if (x < xt[mid]) hi = mid; else lo = mid;
  }
  return lo;
}

// linear_interpolation_module.F90:233-271 (interp_1d_real)
inline double interp_1d(const double *x, const double *f, int n, double xx) {
  int i1 = dintrv(x, n, xx);
  int i2 = i1 + 1;
  double q1 = (xx - x[i1]) / (x[i2] - x[i1]);
  double p1 = 1.0 - q1;
  return p1 * f[i1] + q1 * f[i2];
}

// linear_interpolation_module.F90:297-337 (interp_2d); f is (nx,ny) column-major with leading
// stride `sx` between consecutive x (so a (g,:,:,l) slice of log10k(G,np,nt,nw) can be addressed
// in place: sx = G, sy = G*np).
inline double interp_2d(const double *x, int nx, const double *y, int ny, const double *f,
                        int64_t sx, int64_t sy, double xx, double yy) {
  int ix1 = dintrv(x, nx, xx), ix2 = ix1 + 1;
  int iy1 = dintrv(y, ny, yy), iy2 = iy1 + 1;
  double q1 = (xx - x[ix1]) / (x[ix2] - x[ix1]);
  double q2 = (yy - y[iy1]) / (y[iy2] - y[iy1]);
  double p1 = 1.0 - q1;
  double p2 = 1.0 - q2;
  double fx1 = p1 * f[ix1 * sx + iy1 * sy] + q1 * f[ix2 * sx + iy1 * sy];
  double fx2 = p1 * f[ix1 * sx + iy2 * sy] + q1 * f[ix2 * sx + iy2 * sy];
  return p2 * fx1 + q2 * fx2;
}

// src/clima_eqns.f90:75-80
inline double ten2power(double y) {
  static const double c = std::log(10.0);
  return std::exp(y * c);
}

// src/clima_eqns.f90:64-73
inline double planck_fcn(double nu, double T) {
  return 1.0e3 * ((2.0 * plank * (nu * nu * nu)) / (c_light * c_light)) *
         ((1.0) / (std::exp((plank * nu) / (k_boltz_si * T)) - 1.0));
}

// src/clima_eqns.f90:43-54
inline void weights_to_bins(const double *w, int n, double *bins) {
  bins[0] = 0.0;
  for (int i = 1; i < n + 1; i++) bins[i] = w[i - 1] + bins[i - 1];
}

// futils `is_close` (un-vendored; call sites clima_radtran_types.f90:625-629).  Restated as the
// Fortran-stdlib relative test abs(a-b) <= tol*max(|a|,|b|).  The exact formula is immaterial for
// parity: the doubled grid holds exact copies (src/adiabat/clima_adiabat.f90:736-763).
inline bool is_close(double a, double b, double tol) {
  return std::fabs(a - b) <= tol * std::max(std::fabs(a), std::fabs(b));
}

// futils_mrgrnk `mrgrnk` (un-vendored ORDERPACK 2.0 routine; call site
// clima_radtran_types.f90:840).  Semantics: inds such that x(inds(1)) <= x(inds(2)) <= ...,
// equal keys keep their original relative order (stable bottom-up merge sort on indices).
// `inds` is 0-based here; `tmp` is scratch of size n.
void mrgrnk(const double *x, int n, int *inds, int *tmp) {
  for (int i = 0; i < n; i++) inds[i] = i;
  // order pairs first, as ORDERPACK does
  for (int i = 0; i + 1 < n; i += 2)
    if (x[inds[i]] > x[inds[i + 1]]) std::swap(inds[i], inds[i + 1]);
  int *src = inds, *dst = tmp;
  for (int width = 2; width < n; width *= 2) {
    for (int lo = 0; lo < n; lo += 2 * width) {
      int mid = std::min(lo + width, n), hi = std::min(lo + 2 * width, n);
      int a = lo, b = mid, k = lo;
      while (a < mid && b < hi) {
        if (x[src[b]] < x[src[a]]) dst[k++] = src[b++];  // strict: ties take the earlier run
        else dst[k++] = src[a++];
      }
      while (a < mid) dst[k++] = src[a++];
      while (b < hi) dst[k++] = src[b++];
    }
    std::swap(src, dst);
  }
  if (src != inds) std::memcpy(inds, src, sizeof(int) * n);
}

// futils `rebin` (un-vendored; call site clima_radtran_types.f90:847).  Integral-conserving
// rebin of a piecewise-constant function: new(i) = sum_j old(j)*|old_j ∩ new_i| / |new_i|,
// accumulated in ascending j, each term formed as (b2/b1)*old(j) with b2 the overlap length and
// b1 the new-bin width; an old bin covering the whole new bin contributes old(j) itself.
void rebin(const double *old_bins, const double *old_vals, int n_old, const double *new_bins,
           double *new_vals, int n_new) {
  int l = 0;
  for (int i = 0; i < n_new; i++) {
    new_vals[i] = 0.0;
    const double n0 = new_bins[i], n1 = new_bins[i + 1];
    const double b1 = n1 - n0;
    for (int j = l; j < n_old; j++) {
      const double o0 = old_bins[j], o1 = old_bins[j + 1];
      if (o0 <= n0 && o1 > n0 && o1 < n1) {
        // old bin straddles the left edge of the new bin
        double b2 = o1 - n0;
        new_vals[i] += (b2 / b1) * old_vals[j];
      } else if (o0 <= n0 && o1 >= n1) {
        // old bin covers the new bin
        new_vals[i] += old_vals[j];
        l = j;
        break;
      } else if (o0 > n0 && o1 < n1) {
        // old bin inside the new bin
        double b2 = o1 - o0;
        new_vals[i] += (b2 / b1) * old_vals[j];
      } else if (o0 > n0 && o0 < n1 && o1 >= n1) {
        // old bin straddles the right edge of the new bin
        double b2 = n1 - o0;
        new_vals[i] += (b2 / b1) * old_vals[j];
        l = j;
        break;
      } else if (o0 >= n1) {
        l = j;
        break;
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------
// src/radtran/clima_radtran_twostream.f90:297-316 (tridiag, Thomas, in place)
void tridiag(int n, double *a, double *b, double *c, double *d) {
  c[0] = c[0] / b[0];
  d[0] = d[0] / b[0];
  for (int i = 1; i < n - 1; i++) {
    c[i] = c[i] / (b[i] - a[i] * c[i - 1]);
    d[i] = (d[i] - a[i] * d[i - 1]) / (b[i] - a[i] * c[i - 1]);
  }
  d[n - 1] = (d[n - 1] - a[n - 1] * d[n - 2]) / (b[n - 1] - a[n - 1] * c[n - 2]);
  for (int i = n - 2; i >= 0; i--) d[i] = d[i] - c[i] * d[i + 1];
}

struct TwoStreamWork {
  std::vector<double> tau, w0, gt, gam1, gam2, gam3, gam4, lambda, cap_gam, e1, e2, e3, e4, tauc,
      direct, cp0, cpb, cm0, cmb, A, B, D, E, y1, y2;
  void resize(int nz) {
    for (auto *v : {&tau, &w0, &gt, &gam1, &gam2, &gam3, &gam4, &lambda, &cap_gam, &e1, &e2, &e3,
                    &e4, &cp0, &cpb, &cm0, &cmb, &y1, &y2})
      v->resize(nz);
    tauc.resize(nz + 1);
    direct.resize(nz + 1);
    for (auto *v : {&A, &B, &D, &E}) v->resize(2 * nz);
  }
};

// src/radtran/clima_radtran_twostream.f90:10-154 (two_stream_solar)
void two_stream_solar(int nz, const double *tau_in, const double *w0_in, const double *gt_in,
                      double u0, double Rsfc, double *amean, double *surface_radiance, double *fup,
                      double *fdn, TwoStreamWork &w) {
  const double u1 = 1.0 / std::sqrt(3.0);
  const double Fs_pi = 1.0;
  const double sq3 = std::sqrt(3.0);
  for (int i = 0; i < nz; i++) {
    // :38-40 delta-Eddington scaling
    w.tau[i] = tau_in[i] * (1.0 - w0_in[i] * gt_in[i] * gt_in[i]);
    w.w0[i] = w0_in[i] * (1.0 - gt_in[i] * gt_in[i]) / (1.0 - w0_in[i] * gt_in[i] * gt_in[i]);
    w.gt[i] = gt_in[i] / (1.0 + gt_in[i]);
  }
  for (int i = 0; i < nz; i++) {
    // :43-46
    w.gam1[i] = sq3 * (2.0 - w.w0[i] * (1 + w.gt[i])) / 2.0;
    w.gam2[i] = sq3 * w.w0[i] * (1.0 - w.gt[i]) / 2.0;
    w.gam3[i] = (1.0 - sq3 * w.gt[i] * u0) / 2.0;
    w.gam4[i] = 1.0 - w.gam3[i];
    // :50-51
    w.lambda[i] = std::sqrt(w.gam1[i] * w.gam1[i] - w.gam2[i] * w.gam2[i]);
    w.cap_gam[i] = w.gam2[i] / (w.gam1[i] + w.lambda[i]);
  }
  for (int i = 0; i < nz; i++) {
    // :55-61
    double wrk = std::exp(-w.lambda[i] * w.tau[i]);
    w.e1[i] = 1.0 + w.cap_gam[i] * wrk;
    w.e2[i] = 1.0 - w.cap_gam[i] * wrk;
    w.e3[i] = w.cap_gam[i] + wrk;
    w.e4[i] = w.cap_gam[i] - wrk;
  }
  // :64-67
  w.tauc[0] = 0.0;
  for (int i = 1; i < nz + 1; i++) w.tauc[i] = w.tauc[i - 1] + w.tau[i - 1];
  // :73-87
  w.direct[0] = u0 * Fs_pi;
  for (int i = 0; i < nz; i++) {
    double facp = w.w0[i] * Fs_pi * ((w.gam1[i] - 1.0 / u0) * w.gam3[i] + w.gam4[i] * w.gam2[i]);
    double facm = w.w0[i] * Fs_pi * ((w.gam1[i] + 1.0 / u0) * w.gam4[i] + w.gam2[i] * w.gam3[i]);
    double et0 = std::exp(-w.tauc[i] / u0);
    double etb = et0 * std::exp(-w.tau[i] / u0);
    double denom = w.lambda[i] * w.lambda[i] - 1.0 / (u0 * u0);
    w.direct[i + 1] = u0 * Fs_pi * etb;
    w.cp0[i] = et0 * facp / denom;
    w.cpb[i] = etb * facp / denom;
    w.cm0[i] = et0 * facm / denom;
    w.cmb[i] = etb * facm / denom;
  }
  double Ssfc = Rsfc * w.direct[nz];
  // :93-117 (0-based rows; Fortran row l -> index l-1)
  double *A = w.A.data(), *B = w.B.data(), *D = w.D.data(), *E = w.E.data();
  const double *e1 = w.e1.data(), *e2 = w.e2.data(), *e3 = w.e3.data(), *e4 = w.e4.data();
  const double *cp0 = w.cp0.data(), *cpb = w.cpb.data(), *cm0 = w.cm0.data(), *cmb = w.cmb.data();
  A[0] = 0.0;
  B[0] = e1[0];
  D[0] = -e2[0];
  E[0] = 0.0 - cm0[0];
  for (int i = 0; i < nz - 1; i++) {
    int l = 2 * (i + 1);  // Fortran l = 2*i+1 with 1-based i
    A[l] = e2[i] * e3[i] - e4[i] * e1[i];
    B[l] = e1[i] * e1[i + 1] - e3[i] * e3[i + 1];
    D[l] = e3[i] * e4[i + 1] - e1[i] * e2[i + 1];
    E[l] = e3[i] * (cp0[i + 1] - cpb[i]) + e1[i] * (cmb[i] - cm0[i + 1]);
  }
  for (int i = 0; i < nz - 1; i++) {
    int l = 2 * (i + 1) - 1;  // Fortran l = 2*i
    A[l] = e2[i + 1] * e1[i] - e3[i] * e4[i + 1];
    B[l] = e2[i] * e2[i + 1] - e4[i] * e4[i + 1];
    D[l] = e1[i + 1] * e4[i + 1] - e2[i + 1] * e3[i + 1];
    E[l] = e2[i + 1] * (cp0[i + 1] - cpb[i]) - e4[i + 1] * (cm0[i + 1] - cmb[i]);
  }
  {
    int l = 2 * nz - 1;
    A[l] = e1[nz - 1] - Rsfc * e3[nz - 1];
    B[l] = e2[nz - 1] - Rsfc * e4[nz - 1];
    D[l] = 0.0;
    E[l] = Ssfc - cpb[nz - 1] + Rsfc * cmb[nz - 1];
  }
  tridiag(2 * nz, A, B, D, E);
  for (int i = 0; i < nz; i++) {
    w.y1[i] = E[2 * i];
    w.y2[i] = E[2 * i + 1];
  }
  const double *y1 = w.y1.data(), *y2 = w.y2.data();
  // :135-140
  amean[0] = (1.0 / u1) * (y1[0] * e3[0] - y2[0] * e4[0] + cp0[0]) + w.direct[0] / u0;
  for (int i = 0; i < nz; i++)
    amean[i + 1] =
        (1.0 / u1) * (y1[i] * (e1[i] + e3[i]) + y2[i] * (e2[i] + e4[i]) + cpb[i] + cmb[i]) +
        w.direct[i + 1] / u0;
  // :143-148
  fup[0] = ((y1[0] * e3[0] - y2[0] * e4[0]) + cp0[0]);
  fdn[0] = w.direct[0];
  for (int i = 0; i < nz; i++) {
    fup[i + 1] = (y1[i] * e1[i] + y2[i] * e2[i] + cpb[i]);
    fdn[i + 1] = (y1[i] * e3[i] + y2[i] * e4[i] + cmb[i]) + w.direct[i + 1];
  }
  // :151-152
  int i = nz - 1;
  *surface_radiance = (y1[i] * e3[i] + y2[i] * e4[i] + cmb[i]) / u1 + std::exp(-w.tauc[i + 1] / u0);
}

// src/radtran/clima_radtran_twostream.f90:156-295 (two_stream_ir)
void two_stream_ir(int nz, const double *tau, const double *w0, const double *gt, double emissivity,
                   bool has_hard_surface, double ir_tau_min, const double *bplanck, double *fup,
                   double *fdn, TwoStreamWork &w) {
  const double u1 = 0.5;
  const double norm = 2.0 * pi * u1;
  double Rsfc, Ssfc;
  if (has_hard_surface) Rsfc = 1.0 - emissivity; else Rsfc = 0.0;
  for (int i = 0; i < nz; i++) {
    // :195-201
    w.gam1[i] = 2.0 - w0[i] * (1.0 + gt[i]);
    w.gam2[i] = w0[i] * (1.0 - gt[i]);
    w.lambda[i] = std::sqrt(w.gam1[i] * w.gam1[i] - w.gam2[i] * w.gam2[i]);
    w.cap_gam[i] = w.gam2[i] / (w.gam1[i] + w.lambda[i]);
  }
  for (int i = 0; i < nz; i++) {
    // :205-211
    double wrk = std::exp(-w.lambda[i] * tau[i]);
    w.e1[i] = 1.0 + w.cap_gam[i] * wrk;
    w.e2[i] = 1.0 - w.cap_gam[i] * wrk;
    w.e3[i] = w.cap_gam[i] + wrk;
    w.e4[i] = w.cap_gam[i] - wrk;
  }
  for (int i = 0; i < nz; i++) {
    // :215-234
    double b0n, b1n;
    if (tau[i] <= ir_tau_min) {
      double b_avg = 0.5 * (bplanck[i] + bplanck[i + 1]);
      b0n = b_avg;
      b1n = 0.0;
    } else {
      b0n = bplanck[i];
      b1n = (bplanck[i + 1] - b0n) / tau[i];
    }
    w.cp0[i] = norm * (b0n + b1n * (1.0 / (w.gam1[i] + w.gam2[i])));
    w.cpb[i] = norm * (b0n + b1n * (tau[i] + 1.0 / (w.gam1[i] + w.gam2[i])));
    w.cm0[i] = norm * (b0n + b1n * (-1.0 / (w.gam1[i] + w.gam2[i])));
    w.cmb[i] = norm * (b0n + b1n * (tau[i] - 1.0 / (w.gam1[i] + w.gam2[i])));
  }
  // :236-247
  if (has_hard_surface) {
    Ssfc = emissivity * pi * bplanck[nz];
  } else {
    double b1_bot;
    if (tau[nz - 1] <= ir_tau_min) b1_bot = 0.0;
    else b1_bot = (bplanck[nz] - bplanck[nz - 1]) / tau[nz - 1];
    Ssfc = pi * (bplanck[nz] + u1 * b1_bot);
  }
  double *A = w.A.data(), *B = w.B.data(), *D = w.D.data(), *E = w.E.data();
  const double *e1 = w.e1.data(), *e2 = w.e2.data(), *e3 = w.e3.data(), *e4 = w.e4.data();
  const double *cp0 = w.cp0.data(), *cpb = w.cpb.data(), *cm0 = w.cm0.data(), *cmb = w.cmb.data();
  // :251-275
  A[0] = 0.0;
  B[0] = e1[0];
  D[0] = -e2[0];
  E[0] = 0.0 - cm0[0];
  for (int i = 0; i < nz - 1; i++) {
    int l = 2 * (i + 1);
    A[l] = e2[i] * e3[i] - e4[i] * e1[i];
    B[l] = e1[i] * e1[i + 1] - e3[i] * e3[i + 1];
    D[l] = e3[i] * e4[i + 1] - e1[i] * e2[i + 1];
    E[l] = e3[i] * (cp0[i + 1] - cpb[i]) + e1[i] * (cmb[i] - cm0[i + 1]);
  }
  for (int i = 0; i < nz - 1; i++) {
    int l = 2 * (i + 1) - 1;
    A[l] = e2[i + 1] * e1[i] - e3[i] * e4[i + 1];
    B[l] = e2[i] * e2[i + 1] - e4[i] * e4[i + 1];
    D[l] = e1[i + 1] * e4[i + 1] - e2[i + 1] * e3[i + 1];
    E[l] = e2[i + 1] * (cp0[i + 1] - cpb[i]) - e4[i + 1] * (cm0[i + 1] - cmb[i]);
  }
  {
    int l = 2 * nz - 1;
    A[l] = e1[nz - 1] - Rsfc * e3[nz - 1];
    B[l] = e2[nz - 1] - Rsfc * e4[nz - 1];
    D[l] = 0.0;
    E[l] = Ssfc - cpb[nz - 1] + Rsfc * cmb[nz - 1];
  }
  tridiag(2 * nz, A, B, D, E);
  for (int i = 0; i < nz; i++) {
    w.y1[i] = E[2 * i];
    w.y2[i] = E[2 * i + 1];
  }
  const double *y1 = w.y1.data(), *y2 = w.y2.data();
  // :288-293
  fup[0] = ((y1[0] * e3[0] - y2[0] * e4[0]) + cp0[0]);
  fdn[0] = 0.0;
  for (int i = 0; i < nz; i++) {
    fup[i + 1] = (y1[i] * e1[i] + y2[i] * e2[i] + cpb[i]);
    fdn[i + 1] = (y1[i] * e3[i] + y2[i] * e4[i] + cmb[i]);
  }
}

// ---------------------------------------------------------------------------------------------
// State mirroring `type Radtran` (src/radtran/clima_radtran.f90:31-85) for ONE column.
struct Custom {
  int nP = 0;
  std::vector<double> log10P, dtau_dz, w0, g0;  // (nP), (nP,nw) column-major
};

struct BinWork {  // type OpacityBinWork (clima_radtran_types.f90:165-205)
  std::vector<std::vector<double>> ks;  // nk x (nz*G), (nz,G) column-major
  std::vector<double> cia, pxs, H2O, foreign, w0p, qextp, gtp;
  std::vector<double> tausg, taua, tauc, tausc, w0c, g0c, tausp, tausp_1, taup, taua_1, tau, w0, gt;
  std::vector<double> tau_k, tau_k1, tau_xy, tau_xy1, tau_xy2, wxy1, wxy_e;
  std::vector<int> inds, inds_tmp;
};

struct Oracle {
  clima_rt_tables tb;
  clima_rt_params pr;
  std::vector<double> zen_u, zen_w, albedo, emiss;
  Custom custom;
  bool has_custom = false;
  int nz = 0;
  int nw_ir, nw_sol;
  std::vector<double> freq;         // (nw+1) op%freq
  std::vector<double> wbin, wbin_e, wxy;  // Ksettings
  // opw
  std::vector<double> log10P, log10P_cgs, cols, foreign_col;
  std::vector<char> pair_reuse;
  std::vector<int> ierrs;
  std::vector<BinWork> bins;
  // opr (top-first): tau(nz,G,nw), w0(nz,G,nw), g(nz,nw), tau_band(nz,nw)
  std::vector<double> tau, w0, g, tau_band;
  // wrk_ir / wrk_sol / f_total
  std::vector<double> fup_a_ir, fdn_a_ir, amean_ir, tau_band_ir, fup_n_ir, fdn_n_ir;
  std::vector<double> fup_a_sol, fdn_a_sol, amean_sol, tau_band_sol, fup_n_sol, fdn_n_sol;
  std::vector<double> f_total;
  std::vector<TwoStreamWork> tsw;  // one per thread
  int nthreads = 1;
};

void set_params(Oracle &o, const clima_rt_params *p) {
  o.pr = *p;
  o.zen_u.assign(p->zenith_u, p->zenith_u + p->nzenith);
  o.zen_w.assign(p->zenith_weights, p->zenith_weights + p->nzenith);
  o.albedo.assign(p->surface_albedo, p->surface_albedo + o.nw_sol);
  o.emiss.assign(p->surface_emissivity, p->surface_emissivity + o.nw_ir);
}

void allocate_for_nz(Oracle &o, int nz) {
  if (o.nz == nz) return;
  const clima_rt_tables &tb = o.tb;
  const int G = tb.ngauss, nw = tb.nw;
  o.nz = nz;
  o.log10P.assign(nz, 0.0);
  o.log10P_cgs.assign(nz, 0.0);
  o.cols.assign((size_t)nz * tb.ng, 0.0);
  o.foreign_col.assign(nz, 0.0);
  o.pair_reuse.assign(nz, 0);
  o.ierrs.assign(nw, 0);
  o.bins.resize(nw);
  for (auto &b : o.bins) {
    b.ks.resize(tb.nk);
    for (auto &k : b.ks) k.assign((size_t)nz * G, 0.0);
    b.cia.assign((size_t)nz * std::max(tb.ncia, 1), 0.0);
    b.pxs.assign((size_t)nz * std::max(tb.npxs, 1), 0.0);
    b.H2O.assign(nz, 0.0);
    b.foreign.assign(nz, 0.0);
    b.w0p.assign((size_t)nz * std::max(tb.npart, 1), 0.0);
    b.qextp = b.w0p;
    b.gtp = b.w0p;
    b.tausp_1 = b.w0p;
    for (auto *v : {&b.tausg, &b.taua, &b.tauc, &b.tausc, &b.w0c, &b.g0c, &b.tausp, &b.taup,
                    &b.taua_1, &b.tau, &b.w0, &b.gt})
      v->assign(nz, 0.0);
    b.tau_k.assign((size_t)nz * G, 0.0);
    b.tau_k1.assign(G, 0.0);
    b.tau_xy.assign((size_t)nz * G * G, 0.0);
    b.tau_xy1.assign((size_t)G * G, 0.0);
    b.tau_xy2.assign((size_t)G * G, 0.0);
    b.wxy1.assign((size_t)G * G, 0.0);
    b.wxy_e.assign((size_t)G * G + 1, 0.0);
    b.inds.assign((size_t)G * G, 0);
    b.inds_tmp.assign((size_t)G * G, 0);
  }
  o.tau.assign((size_t)nz * G * nw, 0.0);
  o.w0.assign((size_t)nz * G * nw, 0.0);
  o.g.assign((size_t)nz * nw, 0.0);
  o.tau_band.assign((size_t)nz * nw, 0.0);
  o.fup_a_ir.assign((size_t)(nz + 1) * o.nw_ir, 0.0);
  o.fdn_a_ir = o.fup_a_ir;
  o.amean_ir = o.fup_a_ir;  // allocated and zero, never written (clima_radtran.f90:204-205)
  o.tau_band_ir.assign((size_t)nz * o.nw_ir, 0.0);
  o.fup_n_ir.assign(nz + 1, 0.0);
  o.fdn_n_ir.assign(nz + 1, 0.0);
  o.fup_a_sol.assign((size_t)(nz + 1) * o.nw_sol, 0.0);
  o.fdn_a_sol = o.fup_a_sol;
  o.amean_sol = o.fup_a_sol;
  o.tau_band_sol.assign((size_t)nz * o.nw_sol, 0.0);
  o.fup_n_sol.assign(nz + 1, 0.0);
  o.fdn_n_sol.assign(nz + 1, 0.0);
  o.f_total.assign(nz + 1, 0.0);
  for (auto &t : o.tsw) t.resize(nz);
}

// clima_radtran_types.f90:890-917 (interpolate_Xsection)
void interpolate_Xsection(int dim, int ntemp, const double *temp, const double *data, int l, int nz,
                          const double *T, double *res, const char *pair_reuse) {
  if (dim == 0) {
    for (int j = 0; j < nz; j++) res[j] = data[l];
  } else if (dim == 1) {
    const double T_min = temp[0], T_max = temp[ntemp - 1];
    for (int j = 0; j < nz; j++) {
      if (pair_reuse[j]) {
        res[j] = res[j - 1];
      } else {
        double TT = std::min(std::max(T[j], T_min), T_max);
        double val = interp_1d(temp, data + (size_t)l * ntemp, ntemp, TT);
        res[j] = ten2power(val);
      }
    }
  }
}

// clima_radtran_types.f90:780-888 (k_rorr)
void k_rorr(Oracle &o, int l) {
  const clima_rt_tables &tb = o.tb;
  BinWork &wrk = o.bins[l];
  const int nz = o.nz, G = tb.ngauss, G2 = G * G;
  double *tau_k = wrk.tau_k.data();
  double *tau_xy = wrk.tau_xy.data();
  // :816-819 first k-distribution
  {
    int j1 = tb.k_sp_ind[0] - 1;
    for (int i = 0; i < G; i++)
      for (int j = 0; j < nz; j++)
        tau_k[j + (size_t)i * nz] = wrk.ks[0][j + (size_t)i * nz] * o.cols[j + (size_t)j1 * nz];
  }
  // :823-854 mix the rest
  for (int jj = 1; jj < tb.nk; jj++) {
    int j2 = tb.k_sp_ind[jj] - 1;
    for (int i = 0; i < G; i++)
      for (int j = 0; j < G; j++)
        for (int z = 0; z < nz; z++)
          tau_xy[z + (size_t)(j + i * G) * nz] =
              tau_k[z + (size_t)i * nz] + wrk.ks[jj][z + (size_t)j * nz] * o.cols[z + (size_t)j2 * nz];
    for (int i = 0; i < nz; i++) {
      if (o.pair_reuse[i]) {
        for (int j = 0; j < G; j++) tau_k[i + (size_t)j * nz] = tau_k[i - 1 + (size_t)j * nz];
      } else {
        for (int j = 0; j < G2; j++) wrk.tau_xy1[j] = tau_xy[i + (size_t)j * nz];
        mrgrnk(wrk.tau_xy1.data(), G2, wrk.inds.data(), wrk.inds_tmp.data());
        for (int j = 0; j < G2; j++) {
          wrk.tau_xy2[j] = wrk.tau_xy1[wrk.inds[j]];
          wrk.wxy1[j] = o.wxy[wrk.inds[j]];
        }
        weights_to_bins(wrk.wxy1.data(), G2, wrk.wxy_e.data());
        rebin(wrk.wxy_e.data(), wrk.tau_xy2.data(), G2, o.wbin_e.data(), wrk.tau_k1.data(), G);
        for (int j = 0; j < G; j++) tau_k[i + (size_t)j * nz] = wrk.tau_k1[j];
      }
    }
  }
  // :856-886
  double *tau_band = o.tau_band.data() + (size_t)l * nz;
  for (int j = 0; j < nz; j++) tau_band[j] = 0.0;
  for (int i = 0; i < G; i++) {
    for (int k = 0; k < nz; k++) {
      int n = nz - 1 - k;
      wrk.taua_1[n] = tau_k[k + (size_t)i * nz];
    }
    for (int j = 0; j < nz; j++)
      wrk.tau[j] = wrk.tausg[j] + wrk.taua[j] + wrk.taup[j] + wrk.taua_1[j] + wrk.tauc[j];
    for (int j = 0; j < nz; j++) {
      if (wrk.tau[j] <= tau_min) wrk.w0[j] = 0.0;
      else wrk.w0[j] = std::min(max_w0, (wrk.tausg[j] + wrk.tausp[j] + wrk.tausc[j]) / wrk.tau[j]);
    }
    double *rt = o.tau.data() + ((size_t)l * G + i) * nz;
    double *rw = o.w0.data() + ((size_t)l * G + i) * nz;
    for (int j = 0; j < nz; j++) {
      rt[j] = wrk.tau[j];
      tau_band[j] = tau_band[j] + wrk.tau[j] * o.wbin[i];
      rw[j] = wrk.w0[j];
    }
  }
  double *rg = o.g.data() + (size_t)l * nz;
  for (int j = 0; j < nz; j++) rg[j] = wrk.gt[j];
}

// clima_radtran_types.f90:574-778 (OpticalProperties_compute_opacity)
int compute_opacity_fn(Oracle &o, const double *P, const double *T, const double *densities,
                    const double *dz, const double *pdensities, const double *radii,
                    std::string &err) {
  const clima_rt_tables &tb = o.tb;
  const int nz = o.nz, ng = tb.ng, nw = tb.nw, G = tb.ngauss;
  if (tb.nk == 0) {
    err = "There are no k-distributions, yet there must be some to compute total opacity.";
    return 1;
  }
  // :600-633 prologue
  const double pair_tol = 1.0e-12;
  for (int j = 0; j < nz; j++) {
    o.log10P[j] = std::log10(P[j]);
    o.log10P_cgs[j] = std::log10(P[j] * 1.0e6);
  }
  for (int i = 0; i < ng; i++)
    for (int j = 0; j < nz; j++) o.cols[j + (size_t)i * nz] = densities[j + (size_t)i * nz] * dz[j];
  if (tb.has_cont) {
    for (int j = 0; j < nz; j++) {
      o.foreign_col[j] = 0.0;
      for (int i = 0; i < ng; i++)
        if (i != tb.cont_LH2O - 1) o.foreign_col[j] = o.foreign_col[j] + o.cols[j + (size_t)i * nz];
    }
  }
  std::fill(o.pair_reuse.begin(), o.pair_reuse.end(), 0);
  if (nz % 2 == 0) {
    for (int j = 1; j < nz; j += 2) {  // Fortran j = 2,4,...
      bool r = true;
      r = r && is_close(P[j], P[j - 1], pair_tol);
      r = r && is_close(T[j], T[j - 1], pair_tol);
      for (int i = 0; i < ng; i++)
        r = r && is_close(o.cols[j + (size_t)i * nz], o.cols[j - 1 + (size_t)i * nz], pair_tol);
      if (radii && tb.npart > 0)
        for (int i = 0; i < tb.np; i++)
          r = r && is_close(radii[j + (size_t)i * nz], radii[j - 1 + (size_t)i * nz], pair_tol);
      o.pair_reuse[j] = r;
    }
  }
  std::fill(o.ierrs.begin(), o.ierrs.end(), 0);
  const char *pr = o.pair_reuse.data();

#pragma omp parallel for schedule(dynamic) num_threads(o.nthreads)
  for (int l = 0; l < nw; l++) {
    BinWork &wrk = o.bins[l];
    // :649-662 k-distributions
    for (int i = 0; i < tb.nk; i++) {
      const int npr = tb.k_npress[i], nt = tb.k_ntemp[i];
      const double *lp = tb.k_log10P[i], *tt = tb.k_temp[i];
      const double T_min = tt[0], T_max = tt[nt - 1], P_min = lp[0], P_max = lp[npr - 1];
      for (int k = 0; k < G; k++) {
        const double *f = tb.k_log10k[i] + k + (size_t)l * G * npr * nt;
        double *kk = wrk.ks[i].data() + (size_t)k * nz;
        for (int j = 0; j < nz; j++) {
          if (pr[j]) {
            kk[j] = kk[j - 1];
          } else {
            double TT = std::min(std::max(T[j], T_min), T_max);
            double log10PP = std::min(std::max(o.log10P[j], P_min), P_max);
            double v = interp_2d(lp, npr, tt, nt, f, G, (int64_t)G * npr, log10PP, TT);
            kk[j] = ten2power(v);
          }
        }
      }
    }
    // :665-672 CIA / photolysis
    for (int i = 0; i < tb.ncia; i++)
      interpolate_Xsection(tb.cia_dim[i], tb.cia_ntemp[i], tb.cia_temp[i], tb.cia_data[i], l, nz, T,
                           wrk.cia.data() + (size_t)i * nz, pr);
    for (int i = 0; i < tb.npxs; i++)
      interpolate_Xsection(tb.pxs_dim[i], tb.pxs_ntemp[i], tb.pxs_temp[i], tb.pxs_data[i], l, nz, T,
                           wrk.pxs.data() + (size_t)i * nz, pr);
    // :675-677, 919-945 water continuum
    if (tb.has_cont) {
      const int nt = tb.cont_ntemp;
      const double T_min = tb.cont_temp[0], T_max = tb.cont_temp[nt - 1];
      for (int j = 0; j < nz; j++) {
        if (pr[j]) {
          wrk.H2O[j] = wrk.H2O[j - 1];
          wrk.foreign[j] = wrk.foreign[j - 1];
        } else {
          double TT = std::min(std::max(T[j], T_min), T_max);
          double val = interp_1d(tb.cont_temp, tb.cont_log10_H2O + (size_t)l * nt, nt, TT);
          wrk.H2O[j] = ten2power(val);
          val = interp_1d(tb.cont_temp, tb.cont_log10_foreign + (size_t)l * nt, nt, TT);
          wrk.foreign[j] = ten2power(val);
        }
      }
    }
    // :680-683, 947-983 particles
    for (int i = 0; i < tb.npart; i++) {
      const int nr = tb.part_nrad[i], pind = tb.part_p_ind[i] - 1;
      const double r_min = tb.part_radii[i][0], r_max = tb.part_radii[i][nr - 1];
      int ierr = 0;
      double *w0p = wrk.w0p.data() + (size_t)i * nz, *qe = wrk.qextp.data() + (size_t)i * nz,
             *gp = wrk.gtp.data() + (size_t)i * nz;
      for (int j = 0; j < nz; j++) {
        if (pr[j]) {
          w0p[j] = w0p[j - 1];
          qe[j] = qe[j - 1];
          gp[j] = gp[j - 1];
          continue;
        }
        double rp = radii[j + (size_t)pind * nz];
        if (rp < r_min || rp > r_max) {
          rp = std::min(std::max(rp, r_min), r_max);
          ierr = 1;
        }
        w0p[j] = interp_1d(tb.part_radii[i], tb.part_w0[i] + (size_t)l * nr, nr, rp);
        qe[j] = interp_1d(tb.part_radii[i], tb.part_qext[i] + (size_t)l * nr, nr, rp);
        gp[j] = interp_1d(tb.part_radii[i], tb.part_gt[i] + (size_t)l * nr, nr, rp);
      }
      o.ierrs[l] += ierr;
    }
    // :686-693 Rayleigh
    for (int n = 0; n < nz; n++) wrk.tausg[n] = 0.0;
    for (int i = 0; i < tb.nray; i++) {
      int j = tb.ray_sp_ind[i] - 1;
      for (int k = 0; k < nz; k++) {
        int n = nz - 1 - k;
        wrk.tausg[n] = wrk.tausg[n] + tb.ray_xs[l + (size_t)i * nw] * o.cols[k + (size_t)j * nz];
      }
    }
    // :696-704 CIA
    for (int n = 0; n < nz; n++) wrk.taua[n] = 0.0;
    for (int i = 0; i < tb.ncia; i++) {
      int j = tb.cia_sp_ind[2 * i] - 1, jj = tb.cia_sp_ind[2 * i + 1] - 1;
      for (int k = 0; k < nz; k++) {
        int n = nz - 1 - k;
        wrk.taua[n] = wrk.taua[n] + wrk.cia[k + (size_t)i * nz] * densities[k + (size_t)j * nz] *
                                        densities[k + (size_t)jj * nz] * dz[k];
      }
    }
    // :707-713 photolysis/absorption
    for (int i = 0; i < tb.npxs; i++) {
      int j = tb.pxs_sp_ind[i] - 1;
      for (int k = 0; k < nz; k++) {
        int n = nz - 1 - k;
        wrk.taua[n] = wrk.taua[n] + wrk.pxs[k + (size_t)i * nz] * o.cols[k + (size_t)j * nz];
      }
    }
    // :716-723 continuum
    if (tb.has_cont) {
      int lw = tb.cont_LH2O - 1;
      for (int k = 0; k < nz; k++) {
        int n = nz - 1 - k;
        wrk.taua[n] = wrk.taua[n] +
                      wrk.H2O[k] * densities[k + (size_t)lw * nz] * o.cols[k + (size_t)lw * nz] +
                      wrk.foreign[k] * densities[k + (size_t)lw * nz] * o.foreign_col[k];
      }
    }
    // :726-730, 546-572 custom opacity (flipped to top-first)
    if (!o.has_custom) {
      const double tiny = std::numeric_limits<double>::min();
      for (int n = 0; n < nz; n++) {
        wrk.tauc[n] = tiny;
        wrk.w0c[n] = tiny;
        wrk.g0c[n] = tiny;
      }
    } else {
      const Custom &c = o.custom;
      for (int j = 0; j < nz; j++) {
        int n = nz - 1 - j;
        double t = interp_1d(c.log10P.data(), c.dtau_dz.data() + (size_t)l * c.nP, c.nP, o.log10P_cgs[j]);
        wrk.tauc[n] = t * dz[j];
        wrk.w0c[n] = interp_1d(c.log10P.data(), c.w0.data() + (size_t)l * c.nP, c.nP, o.log10P_cgs[j]);
        wrk.g0c[n] = interp_1d(c.log10P.data(), c.g0.data() + (size_t)l * c.nP, c.nP, o.log10P_cgs[j]);
      }
    }
    for (int n = 0; n < nz; n++) wrk.tausc[n] = wrk.w0c[n] * wrk.tauc[n];
    // :733-744 particles
    for (int n = 0; n < nz; n++) {
      wrk.tausp[n] = 0.0;
      wrk.taup[n] = 0.0;
    }
    for (int i = 0; i < tb.npart; i++) {
      int j = tb.part_p_ind[i] - 1;
      for (int k = 0; k < nz; k++) {
        int n = nz - 1 - k;
        double r = radii[k + (size_t)j * nz];
        double taup_1 = wrk.qextp[k + (size_t)i * nz] * pi * (r * r) * pdensities[k + (size_t)j * nz] * dz[k];
        wrk.taup[n] = wrk.taup[n] + taup_1;
        wrk.tausp_1[n + (size_t)i * nz] = wrk.w0p[k + (size_t)i * nz] * taup_1;
        wrk.tausp[n] = wrk.tausp[n] + wrk.tausp_1[n + (size_t)i * nz];
      }
    }
    // :746-757 asymmetry
    for (int n = 0; n < nz; n++) wrk.gt[n] = 0.0;
    for (int i = 0; i < tb.npart; i++) {
      for (int k = 0; k < nz; k++) {
        int n = nz - 1 - k;
        wrk.gt[n] = wrk.gt[n] + wrk.gtp[k + (size_t)i * nz] * wrk.tausp_1[n + (size_t)i * nz] /
                                    std::max(tau_min, (wrk.tausp[n] + wrk.tausg[n] + wrk.tausc[n]));
      }
    }
    for (int n = 0; n < nz; n++)
      wrk.gt[n] = wrk.gt[n] +
                  wrk.g0c[n] * wrk.tausc[n] / std::max(tau_min, (wrk.tausp[n] + wrk.tausg[n] + wrk.tausc[n]));
    for (int n = 0; n < nz; n++) wrk.gt[n] = std::min(wrk.gt[n], max_gt);
    // :759-760
    k_rorr(o, l);
  }
  for (int l = 0; l < nw; l++)
    if (o.ierrs[l] != 0) {
      err = "Opacity computation failed in one or more wavelength bins.";
      return 1;
    }
  return 0;
}

// src/radtran/clima_radtran_radiate.f90:7-196 (radiate) for one channel
void radiate_channel(Oracle &o, bool is_ir, double diurnal_fac, const double *photons_sol,
                     double photon_scale, const double *zen_u, const double *zen_w, int nzen,
                     double T_surface, const double *T, double *fup_a, double *fdn_a, double *fup_n,
                     double *fdn_n, double *amean, double *tau_band_out) {
  const clima_rt_tables &tb = o.tb;
  const int nz = o.nz, G = tb.ngauss;
  const int ind_start = (is_ir ? tb.ir_ind_start : tb.sol_ind_start) - 1;
  const int ind_end = (is_ir ? tb.ir_ind_end : tb.sol_ind_end) - 1;
  const int nwch = ind_end - ind_start + 1;

#pragma omp parallel for schedule(dynamic) num_threads(o.nthreads)
  for (int l = ind_start; l <= ind_end; l++) {
#ifdef _OPENMP
    TwoStreamWork &tsw = o.tsw[omp_get_thread_num()];
#else
    TwoStreamWork &tsw = o.tsw[0];
#endif
    const int ll = l - ind_start;
    std::vector<double> bpl(nz + 1), fup0(nz + 1), fup1(nz + 1), fup2(nz + 1), fdn0(nz + 1),
        fdn1(nz + 1), fdn2(nz + 1), am0(nz + 1), am1(nz + 1), am2(nz + 1);
    double emissivity = 0.0, albedo = 0.0, surf_rad;
    if (is_ir) {
      // :63-71
      double avg_freq = 0.5 * (o.freq[l] + o.freq[l + 1]);
      bpl[nz] = planck_fcn(avg_freq, T_surface);
      for (int j = 0; j < nz; j++) {
        int n = nz - 1 - j;
        bpl[n] = planck_fcn(avg_freq, T[j]);
      }
      emissivity = o.emiss[ll];
    } else {
      albedo = o.albedo[ll];
    }
    std::fill(fup2.begin(), fup2.end(), 0.0);
    std::fill(fdn2.begin(), fdn2.end(), 0.0);
    std::fill(am2.begin(), am2.end(), 0.0);
    for (int ii = 0; ii < nzen; ii++) {
      std::fill(fup1.begin(), fup1.end(), 0.0);
      std::fill(fdn1.begin(), fdn1.end(), 0.0);
      std::fill(am1.begin(), am1.end(), 0.0);
      for (int i = 0; i < G; i++) {
        const double *tau = o.tau.data() + ((size_t)l * G + i) * nz;
        const double *w0 = o.w0.data() + ((size_t)l * G + i) * nz;
        const double *gt = o.g.data() + (size_t)l * nz;
        if (!is_ir) {
          two_stream_solar(nz, tau, w0, gt, zen_u[ii], albedo, am0.data(), &surf_rad, fup0.data(),
                           fdn0.data(), tsw);
        } else {
          two_stream_ir(nz, tau, w0, gt, emissivity, o.pr.has_hard_surface != 0, o.pr.ir_tau_min,
                        bpl.data(), fup0.data(), fdn0.data(), tsw);
        }
        // :122-126
        for (int n = 0; n < nz + 1; n++) {
          fup1[n] = fup1[n] + fup0[n] * o.wbin[i];
          fdn1[n] = fdn1[n] + fdn0[n] * o.wbin[i];
          if (!is_ir) am1[n] = am1[n] + am0[n] * o.wbin[i];
        }
      }
      // :130-134
      for (int n = 0; n < nz + 1; n++) {
        fup2[n] = fup2[n] + fup1[n] * zen_w[ii];
        fdn2[n] = fdn2[n] + fdn1[n] * zen_w[ii];
        if (!is_ir) am2[n] = am2[n] + am1[n] * zen_w[ii];
      }
    }
    // :140-154 reverse to ground-first
    for (int i = 0; i < nz + 1; i++) {
      int n = nz - i;
      fup_a[i + (size_t)ll * (nz + 1)] = fup2[n];
      fdn_a[i + (size_t)ll * (nz + 1)] = fdn2[n];
      if (!is_ir) amean[i + (size_t)ll * (nz + 1)] = am2[n];
    }
    for (int i = 0; i < nz; i++) {
      int n = nz - 1 - i;
      tau_band_out[i + (size_t)ll * nz] = o.tau_band[n + (size_t)l * nz];
    }
  }
  // :160-192 epilogue.  rtc%freq(ll) == op%freq(ind_start+ll) (types_create.f90:264-266 builds it
  // from the channel's own wavl, which matches op%wavl within 1e-7; synthetic tables use the same
  // array so it is exact here).
  const double *freq = o.freq.data() + ind_start;
  const double *wavl = tb.wavl + ind_start;
  if (!is_ir) {
    for (int l = 0; l < nwch; l++) {
      double ps = photons_sol[l] * photon_scale;  // clima_radtran.f90:302
      for (int i = 0; i < nz + 1; i++) {
        size_t ix = i + (size_t)l * (nz + 1);
        fup_a[ix] = fup_a[ix] * ps * diurnal_fac;
        fdn_a[ix] = fdn_a[ix] * ps * diurnal_fac;
        amean[ix] = amean[ix] * ps * diurnal_fac;
      }
      double avg_freq = 0.5 * (freq[l] + freq[l + 1]);
      double avg_wavl = 1.0e9 * c_light / avg_freq;
      for (int i = 0; i < nz + 1; i++) {
        size_t ix = i + (size_t)l * (nz + 1);
        amean[ix] = amean[ix] * (avg_freq / avg_wavl);
        amean[ix] = amean[ix] * (avg_wavl / (plank * c_light * 1.0e16)) * (wavl[l + 1] - wavl[l]);
      }
    }
  }
  for (int i = 0; i < nz + 1; i++) {
    fup_n[i] = 0.0;
    fdn_n[i] = 0.0;
  }
  for (int l = 0; l < nwch; l++) {
    double dfreq = freq[l] - freq[l + 1];
    for (int i = 0; i < nz + 1; i++) {
      fup_n[i] = fup_n[i] + fup_a[i + (size_t)l * (nz + 1)] * dfreq;
      fdn_n[i] = fdn_n[i] + fdn_a[i + (size_t)l * (nz + 1)] * dfreq;
    }
  }
}

void set_err(char *err, const std::string &s) {
  if (!err) return;
  size_t n = std::min<size_t>(s.size(), CLIMA_RT_ERR_LEN);
  std::memcpy(err, s.data(), n);
  err[n] = 0;
}

// src/radtran/clima_radtran.f90:221-318 (Radtran_radiate) incl. check_inputs (:417-491)
int radtran_radiate(Oracle &o, int nz, int ng, int np, double T_surface, const double *T,
                    const double *P, const double *densities, const double *dz,
                    const double *pdensities, const double *radii, bool compute_solar,
                    bool compute_opacity, std::string &err) {
  if ((pdensities && !radii) || (radii && !pdensities)) {
    err = "Both pdensities and radii must be arguments.";
    return 1;
  }
  if (o.tb.np > 0 && !radii) {
    err = "The model contains particles but \"pdensities\" and \"radii\" are not arguments.";
    return 1;
  }
  if (ng != o.tb.ng) {
    err = "\"densities\" has the wrong input dimension.";
    return 1;
  }
  if (radii && np != o.tb.np) {
    err = "\"pdensities\" has the wrong input dimension.";
    return 1;
  }
  if (nz < 1) {
    err = "\"nz\" can not be less than 1.";
    return 1;
  }
  allocate_for_nz(o, nz);
  if (compute_opacity) {
    if (compute_opacity_fn(o, P, T, densities, dz, pdensities, radii, err)) return 1;
  }
  // :262-283 IR channel: zenith_u=[0], weights=[1], diurnal_fac=0, photons_sol=[0]
  const double zu0[1] = {0.0}, zw1[1] = {1.0}, ps0[1] = {0.0};
  radiate_channel(o, true, 0.0, ps0, 1.0, zu0, zw1, 1, T_surface, T, o.fup_a_ir.data(),
                  o.fdn_a_ir.data(), o.fup_n_ir.data(), o.fdn_n_ir.data(), o.amean_ir.data(),
                  o.tau_band_ir.data());
  if (!compute_solar) {
    // :286-289 stale solar fluxes
    for (int i = 0; i < nz + 1; i++)
      o.f_total[i] = (o.fdn_n_sol[i] - o.fup_n_sol[i]) + (o.fdn_n_ir[i] - o.fup_n_ir[i]);
    return 0;
  }
  // :292-313 solar channel
  radiate_channel(o, false, o.pr.diurnal_fac, o.tb.photons_sol, o.pr.photon_scale_factor,
                  o.zen_u.data(), o.zen_w.data(), (int)o.zen_u.size(), T_surface, T,
                  o.fup_a_sol.data(), o.fdn_a_sol.data(), o.fup_n_sol.data(), o.fdn_n_sol.data(),
                  o.amean_sol.data(), o.tau_band_sol.data());
  // :316
  for (int i = 0; i < nz + 1; i++)
    o.f_total[i] = (o.fdn_n_sol[i] - o.fup_n_sol[i]) + (o.fdn_n_ir[i] - o.fup_n_ir[i]);
  return 0;
}

void copy_out(const Oracle &o, int c, const clima_rt_out *out) {
  const int nz = o.nz;
  const size_t L = nz + 1;
  auto cp = [](double *dst, const std::vector<double> &v, size_t off) {
    if (dst) std::memcpy(dst + off, v.data(), v.size() * sizeof(double));
  };
  cp(out->fup_n_ir, o.fup_n_ir, c * L);
  cp(out->fdn_n_ir, o.fdn_n_ir, c * L);
  cp(out->fup_n_sol, o.fup_n_sol, c * L);
  cp(out->fdn_n_sol, o.fdn_n_sol, c * L);
  cp(out->f_total, o.f_total, c * L);
  // clima_radtran.f90:339-340
  if (out->ISR) out->ISR[c] = (o.fdn_n_sol[nz] - o.fup_n_sol[nz]);
  if (out->OLR) out->OLR[c] = -(o.fdn_n_ir[nz] - o.fup_n_ir[nz]);
  cp(out->fup_a_ir, o.fup_a_ir, c * L * o.nw_ir);
  cp(out->fdn_a_ir, o.fdn_a_ir, c * L * o.nw_ir);
  cp(out->tau_band_ir, o.tau_band_ir, (size_t)c * nz * o.nw_ir);
  cp(out->fup_a_sol, o.fup_a_sol, c * L * o.nw_sol);
  cp(out->fdn_a_sol, o.fdn_a_sol, c * L * o.nw_sol);
  cp(out->amean_sol, o.amean_sol, c * L * o.nw_sol);
  cp(out->tau_band_sol, o.tau_band_sol, (size_t)c * nz * o.nw_sol);
  if (out->status) out->status[c] = 0;
}

Oracle *make_oracle(const clima_rt_tables *tb, const clima_rt_params *pr, int nthreads) {
  Oracle *o = new Oracle;
  o->tb = *tb;
  o->nw_ir = tb->ir_ind_end - tb->ir_ind_start + 1;
  o->nw_sol = tb->sol_ind_end - tb->sol_ind_start + 1;
  set_params(*o, pr);
  const int G = tb->ngauss;
  o->freq.resize(tb->nw + 1);
  // types_create.f90:360-361
  for (int i = 0; i < tb->nw + 1; i++) o->freq[i] = c_light / (tb->wavl[i] * 1.0e-9);
  // create_Ksettings, types_create.f90:191-224
  o->wbin.assign(tb->weights, tb->weights + G);
  o->wbin_e.resize(G + 1);
  weights_to_bins(o->wbin.data(), G, o->wbin_e.data());
  o->wxy.resize((size_t)G * G);
  for (int i = 0; i < G; i++)
    for (int j = 0; j < G; j++) o->wxy[j + (size_t)i * G] = o->wbin[i] * o->wbin[j];
  o->nthreads = std::max(1, nthreads);
  o->tsw.resize(o->nthreads);
  return o;
}

}  // namespace

// =============================================================================================
// C entry points (loaded with ctypes by tests/ and bench.py only)
// =============================================================================================
extern "C" {

void oracle_create(const clima_rt_tables *tables, const clima_rt_params *params,
                   const int32_t *nthreads_bins, void **handle, char *err) {
  set_err(err, "");
  *handle = make_oracle(tables, params, *nthreads_bins);
}

void oracle_destroy(void *handle) { delete static_cast<Oracle *>(handle); }

void oracle_set_params(void *handle, const clima_rt_params *params, char *err) {
  set_err(err, "");
  set_params(*static_cast<Oracle *>(handle), params);
}

void oracle_set_custom_optical_properties(void *handle, const int32_t *nP, const double *log10P,
                                          const double *dtau_dz, const double *w0, const double *g0,
                                          char *err) {
  set_err(err, "");
  Oracle &o = *static_cast<Oracle *>(handle);
  const size_t n = (size_t)(*nP) * o.tb.nw;
  o.custom.nP = *nP;
  o.custom.log10P.assign(log10P, log10P + *nP);
  o.custom.dtau_dz.assign(dtau_dz, dtau_dz + n);
  o.custom.w0.assign(w0, w0 + n);
  o.custom.g0.assign(g0, g0 + n);
  o.has_custom = true;
}

void oracle_unset_custom_optical_properties(void *handle) {
  static_cast<Oracle *>(handle)->has_custom = false;
}

// Radtran%radiate + TOA_fluxes for ONE column (state kept in the handle like the Fortran type).
void oracle_radiate(void *handle, const int32_t *nz, const int32_t *ng, const int32_t *np,
                    const double *T_surface, const double *T, const double *P,
                    const double *densities, const double *dz, const double *pdensities,
                    const double *radii, const int32_t *compute_solar,
                    const int32_t *compute_opacity, const clima_rt_out *out, char *err) {
  set_err(err, "");
  Oracle &o = *static_cast<Oracle *>(handle);
  std::string e;
  if (radtran_radiate(o, *nz, *ng, *np, *T_surface, T, P, densities, dz, pdensities, radii,
                      *compute_solar != 0, *compute_opacity != 0, e)) {
    set_err(err, e);
    return;
  }
  copy_out(o, 0, out);
}

// The batched entry: columns are independent; parallelised over columns with OpenMP (the best
// case for a CPU; the reference itself only threads over wavelength bins inside one column).
// With compute_opacity=0 every column reuses the opacity of the handle (frozen-opacity Jacobian
// mode, src/adiabat/clima_adiabat_solve.f90:798-820).
void oracle_radiate_batch(void *handle, const int32_t *ncol, const int32_t *nz, const int32_t *ng,
                          const int32_t *np, const double *T_surface, const double *T,
                          const double *P, const double *densities, const double *dz,
                          const double *pdensities, const double *radii,
                          const int32_t *compute_solar, const int32_t *compute_opacity,
                          const int32_t *nthreads_cols, const clima_rt_out *out, char *err) {
  set_err(err, "");
  Oracle &base = *static_cast<Oracle *>(handle);
  const int NZ = *nz, NG = *ng, NP = *np, NC = *ncol;
  const int nth = std::max(1, *nthreads_cols);
  std::vector<std::string> errs(nth);
  allocate_for_nz(base, NZ);
#pragma omp parallel num_threads(nth)
  {
#ifdef _OPENMP
    const int tid = omp_get_thread_num();
#else
    const int tid = 0;
#endif
    Oracle local = base;  // private work arrays (and a copy of the frozen opacity / stale solar)
    local.nthreads = 1;
    local.tsw.assign(1, TwoStreamWork());
    local.tsw[0].resize(NZ);
#pragma omp for schedule(dynamic)
    for (int c = 0; c < NC; c++) {
      std::string e;
      const double *pd = pdensities ? pdensities + (size_t)c * NZ * NP : nullptr;
      const double *rd = radii ? radii + (size_t)c * NZ * NP : nullptr;
      if (radtran_radiate(local, NZ, NG, NP, T_surface[c], T + (size_t)c * NZ, P + (size_t)c * NZ,
                          densities + (size_t)c * NZ * NG, dz + (size_t)c * NZ, pd, rd,
                          *compute_solar != 0, *compute_opacity != 0, e)) {
        if (errs[tid].empty()) errs[tid] = e;
        continue;
      }
      copy_out(local, c, out);
    }
  }
  for (auto &e : errs)
    if (!e.empty()) {
      set_err(err, e);
      return;
    }
}

// OpticalPropertiesResult of the handle's last compute_opacity (top-first, Fortran layout).
void oracle_get_opacity(void *handle, double *tau, double *w0, double *g, double *tau_band) {
  Oracle &o = *static_cast<Oracle *>(handle);
  std::memcpy(tau, o.tau.data(), o.tau.size() * sizeof(double));
  std::memcpy(w0, o.w0.data(), o.w0.size() * sizeof(double));
  std::memcpy(g, o.g.data(), o.g.size() * sizeof(double));
  std::memcpy(tau_band, o.tau_band.data(), o.tau_band.size() * sizeof(double));
}

// ---- unit-level exports used by tests --------------------------------------------------------
void oracle_dintrv(const double *xt, const int32_t *n, const double *x, int32_t *ileft) {
  *ileft = dintrv(xt, *n, *x) + 1;  // 1-based like the Fortran
}
void oracle_interp_1d(const double *x, const double *f, const int32_t *n, const double *xx,
                      double *out) {
  *out = interp_1d(x, f, *n, *xx);
}
void oracle_interp_2d(const double *x, const int32_t *nx, const double *y, const int32_t *ny,
                      const double *f, const double *xx, const double *yy, double *out) {
  *out = interp_2d(x, *nx, y, *ny, f, 1, *nx, *xx, *yy);
}
void oracle_planck(const double *nu, const double *T, double *B) { *B = planck_fcn(*nu, *T); }
void oracle_ten2power(const double *y, double *r) { *r = ten2power(*y); }
void oracle_weights_to_bins(const double *w, const int32_t *n, double *bins) {
  weights_to_bins(w, *n, bins);
}
void oracle_mrgrnk(const double *x, const int32_t *n, int32_t *inds) {
  std::vector<int> a(*n), t(*n);
  mrgrnk(x, *n, a.data(), t.data());
  for (int i = 0; i < *n; i++) inds[i] = a[i] + 1;  // 1-based
}
void oracle_rebin(const double *old_bins, const double *old_vals, const int32_t *n_old,
                  const double *new_bins, double *new_vals, const int32_t *n_new) {
  rebin(old_bins, old_vals, *n_old, new_bins, new_vals, *n_new);
}
void oracle_tridiag(const int32_t *n, double *a, double *b, double *c, double *d) {
  tridiag(*n, a, b, c, d);
}
void oracle_two_stream_ir(const int32_t *nz, const double *tau, const double *w0, const double *gt,
                          const double *emissivity, const int32_t *has_hard_surface,
                          const double *tau_min_ir, const double *bplanck, double *fup, double *fdn) {
  TwoStreamWork w;
  w.resize(*nz);
  two_stream_ir(*nz, tau, w0, gt, *emissivity, *has_hard_surface != 0, *tau_min_ir, bplanck, fup,
                fdn, w);
}
void oracle_two_stream_solar(const int32_t *nz, const double *tau, const double *w0,
                             const double *gt, const double *u0, const double *Rsfc, double *amean,
                             double *surface_radiance, double *fup, double *fdn) {
  TwoStreamWork w;
  w.resize(*nz);
  two_stream_solar(*nz, tau, w0, gt, *u0, *Rsfc, amean, surface_radiance, fup, fdn, w);
}
int32_t oracle_max_threads() {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

}  // extern "C"
