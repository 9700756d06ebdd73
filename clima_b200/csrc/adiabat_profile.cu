// adiabat_profile.cu -- row N4 of SURVEY.md section 8(f): `AdiabatClimate%make_profile` for a batch of columns.
//
// One thread per column.  Each thread integrates the multispecies pseudo-adiabat (T, z)(P) from the surface
// to P_top with DOP853 (dense output, events for "a dry gas reaches saturation" and "T reaches T_trop" located
// with Brent's zeroin), samples it on the log-spaced pressure grid, joins the isothermal stratosphere, then
// integrates the hydrostatic altitude again on the model grid (`compute_altitude`) and writes the column on the
// radiative grid (`copy_atm_to_radiative_grid`) straight into the arrays the RT kernels read.
//
// Reference: src/adiabat/clima_adiabat_general.f90:89-669 (make_profile, integrate, solout_dop, find_root,
// root_fcn, update_f_i_dry, mixing_ratios, right_hand_side, altitude_vs_P), src/adiabat/clima_adiabat.f90:401-472
// and :729-773, src/adiabat/clima_adiabat_altitude.f90:16-243, src/clima_useful.f90:97-236,
// src/clima_saturationdata.f90:56-166, src/clima_eqns.f90:82-133,201-211.  The integrator (jacobwilliams/dop853
// 1.2.0) and the root finder (futils 0.1.14 brent_class) are dependencies that are not in the reference tree:
// both are restated from the published algorithms (Hairer, Norsett, Wanner II.5 / dop853.f; Brent 1973).
// The step sequence of a column depends only on that column, so a thread's result does not depend on the batch.
// Lanes of a warp sit at different places of the integrator most of the time, so the code is kept SMALL (the
// right-hand side, the stages, the event search and the dense output are __noinline__, loops over stages are not
// unrolled): fully inlined it was 429 KB of SASS and the warps stalled on instruction fetch.
#include <cfloat>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <mutex>

#include "dop853_tables.h"
#include "rt_kernels.h"

#ifndef ADIABAT_NOINLINE
#define ADIABAT_NOINLINE __noinline__
#endif

#ifndef NI_MATH
#define NI_MATH ADIABAT_NOINLINE
#endif
#ifndef NI_RHS
#define NI_RHS ADIABAT_NOINLINE
#endif
#ifndef NI_STAGE
#define NI_STAGE ADIABAT_NOINLINE
#endif
#ifndef NI_SOL
#define NI_SOL ADIABAT_NOINLINE
#endif
#ifndef NI_MISC
#define NI_MISC ADIABAT_NOINLINE
#endif

namespace adiabat {

constexpr int MAXG = ADIABAT_MAXG;
constexpr double RGAS = 8.31446261815324e7;  // clima_const.f90:9
constexpr double K_BOLTZ = 1.380649e-16;     // clima_const.f90:10
constexpr double G_GRAV = 6.67430e-11;       // clima_const.f90:12
constexpr double N_AVO = 6.02214076e23;      // clima_const.f90:15

__device__ __forceinline__ double gravity(double radius, double mass, double z) {  // clima_eqns.f90:201-211
  const double r = (radius + z) / 1.0e2;
  return G_GRAV * (mass / 1.0e3) / (r * r) * 1.0e2;
}

// one copy of each libm routine in the code (see the note on code size above)
__device__ NI_MATH double nlog(double x) { return log(x); }
__device__ NI_MATH double nexp(double x) { return exp(x); }
__device__ NI_MATH double nlog10(double x) { return log10(x); }
__device__ NI_MATH double npow(double x, double y) { return pow(x, y); }

__device__ __forceinline__ double integral_fcn(double A, double B, double T) { return -A / T + B * nlog(T); }

// SaturationData%sat_pressure (clima_saturationdata.f90:97-155); s = mu, T_ref, P_ref, T_triple, T_critical,
// a_v, b_v, a_s, b_s, a_c, b_c, then the temperature-independent terms of the three branches (filled by
// clima_rt_adiabat_set in the order the reference adds them): [11] I_v(T_ref), [12] I_v(T_critical) - I_v(T_ref),
// [13] I_c(T_critical), [14] I_v(T_triple) - I_v(T_ref), [15] I_s(T_triple), with I_x(T) = -a_x / T + b_x log T.
__device__ NI_MISC double sat_pressure(const double *s, double T) {
  double tmp;
  if (T >= s[4])
    tmp = s[12] + (integral_fcn(s[9], s[10], T) - s[13]);
  else if (T > s[3])
    tmp = integral_fcn(s[5], s[6], T) - s[11];
  else
    tmp = s[14] + (integral_fcn(s[7], s[8], T) - s[15]);
  return s[2] * nexp((s[0] / RGAS) * tmp);
}

__device__ __forceinline__ double latent_heat(const double *s, double T) {  // :56-91 [erg/g]
  if (T >= s[4]) return s[9] + s[10] * T;
  if (T > s[3]) return s[5] + s[6] * T;
  return s[7] + s[8] * T;
}

// heat_capacity_eval (clima_eqns.f90:105-133) [J/(mol K)]; false when T is outside every range.  The powers of
// T (Shomate: of T / 1000) are the same for every gas and are computed once per right-hand side.
struct TPowers {
  double T, TT, TT2, TT3, rTT2, T2, T3, T4, rT, rT2;
  __device__ explicit TPowers(double T_) : T(T_) {
    TT = T / 1000.0;
    TT2 = TT * TT;
    TT3 = TT * TT * TT;
    rTT2 = 1.0 / TT2;
    T2 = T * T;
    T3 = T2 * T;
    T4 = T2 * T2;
    rT = 1.0 / T;
    rT2 = 1.0 / T2;
  }
};
__device__ bool heat_capacity(const AdiabatData &a, int i, const TPowers &t, double &cp) {
  for (int k = 0; k < a.ntemps[i]; k++) {
    if (t.T >= a.temps[i][k] && t.T < a.temps[i][k + 1]) {
      const double *c = a.coef[i][k];
      if (a.model[i] == 1)
        cp = c[0] + c[1] * t.TT + c[2] * t.TT2 + c[3] * t.TT3 + c[4] * t.rTT2;
      else
        cp = 8.31446261815324 *
             (c[0] * t.rT2 + c[1] * t.rT + c[2] + c[3] * t.T + c[4] * t.T2 + c[5] * t.T3 + c[6] * t.T4);
      return true;
    }
  }
  return false;
}

// AdiabatProfileData (:18-71): what the procedures below share
struct Col {
  const AdiabatData *a;
  int ng;
  double T_trop;           // of this column (AdiabatOut::T_trop_in) or the object's
  signed char cond[MAXG];  // 1 = CondensingSpeciesType, 0 = DrySpeciesType
  double f_i_dry[MAXG];
  int err;
  int nfcn;
};

__device__ void update_f_i_dry(Col &d, double P, const double *f_layer) {  // :529-546
  double P_dry = 0.0;
  for (int i = 0; i < d.ng; i++)
    if (!d.cond[i]) P_dry += f_layer[i] * P;
  for (int i = 0; i < d.ng; i++) d.f_i_dry[i] = (f_layer[i] * P) / P_dry;
}

// returns f_dry (by value: with the function out of line, an unused reference result was given the address of f[0])
__device__ NI_MISC double mixing_ratios(const Col &d, double P, double T, double *f) {  // :548-574
  const AdiabatData &a = *d.a;
  double f_moist = 0.0;
  for (int i = 0; i < d.ng; i++)
    if (d.cond[i]) {
      f[i] = fmin(a.RH[i] * sat_pressure(a.sat[i], T) / P, 1.0);
      f_moist += f[i];
    }
  const double f_dry = fmax(1.0 - f_moist, 1.0e-40);
  for (int i = 0; i < d.ng; i++)
    if (!d.cond[i]) f[i] = f_dry * d.f_i_dry[i];
  return f_dry;
}

// one component of root_fcn (:483-513): 0 = tropopause, i + 1 = gas i reaches saturation
__device__ NI_MISC double root_component(const Col &d, double P, double T, int ind) {
  if (ind == 0) return T - d.T_trop;
  const int i = ind - 1;
  if (d.cond[i]) return 1.0;
  if (!d.a->has_sat[i]) return DBL_MAX;
  double f[MAXG];
  mixing_ratios(d, P, T, f);
  return d.a->RH[i] * sat_pressure(d.a->sat[i], T) - f[i] * P;
}

// right_hand_side (:576-656): Eq. 1 of Graham et al. (2021) without condensate retention, and hydrostatics
struct ProfileRhs {
  Col *d;
  __device__ NI_RHS void operator()(double P, const double *u, double *du) const {
    Col &c = *d;
    const AdiabatData &a = *c.a;
    const double T = u[0], z = u[1];
    const double Rgas_si = RGAS / 1.0e7;
    double f[MAXG];
    const double f_dry = mixing_ratios(c, P, T, f);
    double mubar = 0.0;
    for (int i = 0; i < c.ng; i++) mubar += f[i] * a.mass[i];
    double cp_dry = DBL_MIN, first = 0.0, second = 0.0;
    const TPowers tp(T);
    const double rRT = 1.0 / (Rgas_si * T);
    for (int i = 0; i < c.ng; i++) {
      double cp;
      if (!heat_capacity(a, i, tp, cp)) {
        c.err = 6;
        du[0] = du[1] = 0.0;
        return;
      }
      if (c.cond[i]) {
        const double L = latent_heat(a.sat[i], T) * a.mass[i] * 1.0e-7;  // J/mol
        const double beta = L * rRT;
        first += f[i] * (cp - Rgas_si * beta + Rgas_si * (beta * beta));
        second += beta * f[i];
      } else {
        cp_dry += c.f_i_dry[i] * cp;
      }
    }
    const double dlnT_dlnP = 1.0 / (f_dry * ((cp_dry * f_dry + first) / (Rgas_si * (f_dry + second))) + second);
    du[0] = dlnT_dlnP * (T / P);
    const double grav = gravity(a.planet_radius, a.planet_mass, z);
    du[1] = -(RGAS * T) / (grav * P * mubar);
    c.nfcn++;
  }
};

// ---------------------------------------------------------------------------------------------------------
// DOP853: the driver loop of dop853.f (step-size rule safe 0.9, fac1 0.333, fac2 6, beta 0; hinit; the error
// is the 5th-order estimate damped by the 3rd-order one); dense output contd8.  The stiffness test (every 1000
// accepted steps) is left out: these integrations take < 200 steps, and nmax stops a runaway one.
// ---------------------------------------------------------------------------------------------------------
template <int N>
struct Dop {
  double K[16][N];
  double F[7][N];
  double yold[N];
  double xold, hout;
  __device__ NI_SOL double contd8(int i, double xq) const {
    const double s = (xq - xold) / hout, s1 = 1.0 - s;
    const double conpar = F[3][i] + s * (F[4][i] + s1 * (F[5][i] + s * F[6][i]));
    return yold[i] + s * (F[0][i] + s1 * (F[1][i] + s * (F[2][i] + s1 * conpar)));
  }
};

template <int N, class Rhs>
__device__ NI_STAGE void dop_stage(Dop<N> &d, int s, double x, double h, const double *y, const Rhs &rhs) {
  double y1[N];
#pragma unroll
  for (int i = 0; i < N; i++) {
    double acc = 0.0;
    for (int j = 0; j < s; j++) acc += dop853::A[s][j] * d.K[j][i];
    y1[i] = y[i] + h * acc;
  }
  rhs(x + dop853::C[s] * h, y1, d.K[s]);
}

// -> idid: 1 reached xend, 2 stopped by solout, -2 more than nmax steps, -3 step size too small
template <int N, class Rhs, class Solout>
__device__ int dop_integrate(Dop<N> &d, double &x, double *y, double xend, double rtol, double atol,
                             const Rhs &rhs, Solout &solout) {
  const double safe = 0.9, fac1 = 0.333, fac2 = 6.0, uround = 2.220446049250313e-16;
  const int nmax = 100000;
  double facold = 1.0e-4;
  const double expo1 = 1.0 / 8.0, facc1 = 1.0 / fac1, facc2 = 1.0 / fac2;
  const double posneg = copysign(1.0, xend - x);
  const double hmax = fabs(xend - x);
  bool last = false, reject = false;
  int nstep = 0, naccpt = 0;
  rhs(x, y, d.K[0]);
  double h;
  {  // hinit
    double dnf = 0.0, dny = 0.0, sk[N], y1[N], f1[N];
#pragma unroll
    for (int i = 0; i < N; i++) {
      sk[i] = atol + rtol * fabs(y[i]);
      dnf += (d.K[0][i] / sk[i]) * (d.K[0][i] / sk[i]);
      dny += (y[i] / sk[i]) * (y[i] / sk[i]);
    }
    h = (dnf <= 1.0e-10 || dny <= 1.0e-10) ? 1.0e-6 : sqrt(dny / dnf) * 0.01;
    h = copysign(fmin(h, hmax), posneg);
#pragma unroll
    for (int i = 0; i < N; i++) y1[i] = y[i] + h * d.K[0][i];
    rhs(x + h, y1, f1);
    double der2 = 0.0;
#pragma unroll
    for (int i = 0; i < N; i++) der2 += ((f1[i] - d.K[0][i]) / sk[i]) * ((f1[i] - d.K[0][i]) / sk[i]);
    der2 = sqrt(der2) / h;
    const double der12 = fmax(fabs(der2), sqrt(dnf));
    const double h1 = der12 <= 1.0e-15 ? fmax(1.0e-6, fabs(h) * 1.0e-3) : npow(0.01 / der12, 1.0 / 8.0);
    h = copysign(fmin(fmin(100.0 * fabs(h), h1), hmax), posneg);
  }
  d.xold = x;
  d.hout = 1.0;
  if (solout(d, naccpt + 1, x, x, y) < 0) return 2;
  for (;;) {
    if (nstep > nmax) return -2;
    if (0.1 * fabs(h) <= fabs(x) * uround) return -3;
    if ((x + 1.01 * h - xend) * posneg > 0.0) {
      h = xend - x;
      last = true;
    }
    nstep++;
#pragma unroll 1
    for (int s = 1; s < 12; s++) dop_stage<N>(d, s, x, h, y, rhs);
    const double xph = x + h;
    double ynew[N], err = 0.0, err2 = 0.0;
#pragma unroll
    for (int i = 0; i < N; i++) {
      double k4 = 0.0, e5 = 0.0, e3 = 0.0;
      for (int j = 0; j < 12; j++) {
        k4 += dop853::B[j] * d.K[j][i];
        e5 += dop853::E5[j] * d.K[j][i];
        e3 += dop853::E3[j] * d.K[j][i];
      }
      ynew[i] = y[i] + h * k4;
      const double sk = atol + rtol * fmax(fabs(y[i]), fabs(ynew[i]));
      err += (e5 / sk) * (e5 / sk);
      err2 += (e3 / sk) * (e3 / sk);
    }
    double deno = err + 0.01 * err2;
    if (deno <= 0.0) deno = 1.0;
    err = fabs(h) * err * sqrt(1.0 / (N * deno));
    const double fac11 = npow(err, expo1);
    const double fac = fmax(facc2, fmin(facc1, fac11 / safe));
    double hnew = h / fac;
    if (err <= 1.0) {
      facold = fmax(err, 1.0e-4);
      naccpt++;
      rhs(xph, ynew, d.K[12]);
#pragma unroll 1
      for (int s = 13; s < 16; s++) dop_stage<N>(d, s, x, h, y, rhs);
#pragma unroll
      for (int i = 0; i < N; i++) {
        const double dy = ynew[i] - y[i];
        d.F[0][i] = dy;
        d.F[1][i] = h * d.K[0][i] - dy;
        d.F[2][i] = 2.0 * dy - h * (d.K[12][i] + d.K[0][i]);
        for (int r = 0; r < 4; r++) {
          double acc = 0.0;
          for (int j = 0; j < 16; j++) acc += dop853::D[r][j] * d.K[j][i];
          d.F[3 + r][i] = h * acc;
        }
        d.yold[i] = y[i];
        d.K[0][i] = d.K[12][i];
        y[i] = ynew[i];
      }
      d.hout = h;
      d.xold = x;
      x = xph;
      if (solout(d, naccpt + 1, d.xold, x, y) < 0) return 2;
      if (last) return 1;
      if (fabs(hnew) > hmax) hnew = posneg * hmax;
      if (reject) hnew = posneg * fmin(fabs(hnew), fabs(h));
      reject = false;
    } else {
      hnew = h / fmin(facc1, fac11 / safe);
      reject = true;
      last = false;
    }
    h = hnew;
  }
  (void)facold;
}

// Brent's zeroin (futils brent_class%find_zero); info -1 when f(ax) and f(bx) have one sign
template <class Fn>
__device__ NI_MISC double zeroin(const Fn &f, double ax, double bx, double tol, int &info) {
  const double eps = 2.220446049250313e-16;
  double a = ax, b = bx, fa = f(a), fb = f(b);
  info = 0;
  if (fa * fb > 0.0) {
    info = -1;
    return b;
  }
  double c = a, fc = fa, dd = b - a, e = dd;
  for (;;) {
    if ((fb > 0.0 && fc > 0.0) || (fb < 0.0 && fc < 0.0)) {
      c = a;
      fc = fa;
      dd = e = b - a;
    }
    if (fabs(fc) < fabs(fb)) {
      a = b; b = c; c = a;
      fa = fb; fb = fc; fc = fa;
    }
    const double tol1 = 2.0 * eps * fabs(b) + 0.5 * tol;
    const double xm = 0.5 * (c - b);
    if (fabs(xm) <= tol1 || fb == 0.0) return b;
    if (fabs(e) >= tol1 && fabs(fa) > fabs(fb)) {
      double p, q, s = fb / fa;
      if (a == c) {
        p = 2.0 * xm * s;
        q = 1.0 - s;
      } else {
        q = fa / fc;
        const double r = fb / fc;
        p = s * (2.0 * xm * q * (q - r) - (b - a) * (r - 1.0));
        q = (q - 1.0) * (r - 1.0) * (s - 1.0);
      }
      if (p > 0.0) q = -q; else p = -p;
      if (2.0 * p < fmin(3.0 * xm * q - fabs(tol1 * q), fabs(e * q))) {
        e = dd;
        dd = p / q;
      } else {
        dd = e = xm;
      }
    } else {
      dd = e = xm;
    }
    a = b;
    fa = fb;
    b += fabs(dd) > tol1 ? dd : copysign(tol1, xm);
    fb = f(b);
  }
}

// ---------------------------------------------------------------------------------------------------------
// solout_dop (:355-449): event search on the step just accepted, then the samples on the pressure grid
// ---------------------------------------------------------------------------------------------------------
struct ProfileSolout {
  Col *col;
  // the log-spaced edge grid P_e(0 .. ne-1) (:256-260), generated on the fly
  int ne, j;
  double l0, range, P_surf, Pj;
  double gout_old[MAXG + 1];
  int reason;  // 0 reached P_top, 1 tropopause, 2 a gas reached saturation
  int ind_sat;
  double P_root, u_root[2];
  bool z_negative;
  // model-grid outputs of this column: layer i is edge 2 i + 1
  double *T_m, *f_m;  // T_m[i], f_m[i + nz * g]
  int nz;

  __device__ NI_SOL double P_edge(int k) const {
    if (k == 0) return P_surf;
    if (k == ne - 1) return col->a->P_top;
    return npow(10.0, l0 + range * k / (ne - 1));
  }
  __device__ void sample(int k, double T, double z, const double *f) {
    if (z < 0.0 && k < ne - 1) z_negative = true;
    if (k & 1) {
      const int i = k >> 1;
      T_m[i] = T;
      for (int g = 0; g < col->ng; g++) f_m[i + (size_t)nz * g] = f[g];
    }
  }
  __device__ NI_SOL int operator()(const Dop<2> &dop, int nr, double P_old, double P_cur, const double *y) {
    Col &d = *col;
    if (d.err) return -10;
    int irtrn = 1;
    double gout[MAXG + 1];
    for (int i = 0; i <= d.ng; i++) gout[i] = root_component(d, P_cur, y[0], i);
    if (nr == 1)
      for (int i = 0; i <= d.ng; i++) gout_old[i] = gout[i];
    double best = 0.0;
    int ind_root = -1;
    for (int i = 0; i <= d.ng; i++) {
      if ((gout_old[i] < 0.0 && gout[i] > 0.0) || (gout_old[i] > 0.0 && gout[i] < 0.0)) {
        int info;
        const double r = zeroin([&](double x) { return root_component(d, x, dop.contd8(0, x), i); }, P_old,
                                P_cur, 1.0e-8, info);
        if (info != 0) {
          d.err = 5;
          return -10;
        }
        if (ind_root < 0 || r > best) {  // the highest-pressure root, the first one on a tie (maxloc)
          best = r;
          ind_root = i;
        }
      }
    }
    if (ind_root >= 0) {
      irtrn = -1;
      P_root = best;
      u_root[0] = dop.contd8(0, P_root);
      u_root[1] = dop.contd8(1, P_root);
      if (ind_root == 0) {
        reason = 1;
      } else {
        reason = 2;
        ind_sat = ind_root - 1;
      }
    }
    for (int i = 0; i <= d.ng; i++) gout_old[i] = gout[i];
    if (j < ne) {
      const double PP = irtrn == -1 ? P_root : P_cur;
      if (Pj <= P_old && Pj >= PP) {
        while (Pj >= PP) {
          const double T = dop.contd8(0, Pj), z = dop.contd8(1, Pj);
          double f[MAXG];
          mixing_ratios(d, Pj, T, f);
          sample(j, T, z, f);
          j++;
          if (j >= ne) break;
          Pj = P_edge(j);
        }
      }
    }
    return irtrn;
  }
};

__device__ double altitude_vs_P(double P, double T, double mubar, double P0, double z0, double planet_mass,
                                double planet_radius) {  // :659-669
  const double G_grav_cgs = 6.67e-8;
  return 1.0 / ((N_AVO * K_BOLTZ * T) / (G_grav_cgs * planet_mass * mubar) * nlog(P / P0) + 1.0 / (planet_radius + z0)) -
         planet_radius;
}

// ---------------------------------------------------------------------------------------------------------
// compute_altitude (clima_adiabat_altitude.f90, reference_pressure <= 0): hydrostatics over T and mubar
// interpolated linearly in log10 P between the surface and the layer centres
// ---------------------------------------------------------------------------------------------------------
struct AltitudeRhs {
  // descending in pressure (ground first): lp[0 .. nz], Tg, mg of this column
  const double *lp, *Tg, *mg;
  int nz;
  double planet_mass, planet_radius;
  // linear_interp_1d%evaluate on the ascending grid xg(a) = lp[nz - a]: the bracketing interval, the end
  // intervals outside the grid
  __device__ void interp(double x, double &T, double &m) const {
    const int n = nz + 1;
    const double x0 = lp[nz], x1 = lp[0];
    int a = (int)floor((x - x0) / (x1 - x0) * (n - 1));
    a = min(max(a, 0), n - 2);
    while (a > 0 && x < lp[nz - a]) a--;
    while (a < n - 2 && x >= lp[nz - a - 1]) a++;
    const double xa = lp[nz - a], xb = lp[nz - a - 1];
    const double q = (x - xa) / (xb - xa), p = 1.0 - q;
    T = p * Tg[nz - a] + q * Tg[nz - a - 1];
    m = p * mg[nz - a] + q * mg[nz - a - 1];
  }
  __device__ NI_RHS void operator()(double P, const double *u, double *du) const {
    double T, m;
    interp(nlog10(P), T, m);
    du[0] = -(RGAS * T) / (gravity(planet_radius, planet_mass, u[0]) * P * m);
  }
};

// solout_dop853_solveivp (clima_useful.f90:216-236): dense-output samples at the requested pressures
struct AltitudeSolout {
  const double *x_eval;  // first requested pressure; the j-th is x_eval[j * step]
  double *y_eval;
  int m, j, step;
  double xtol;
  __device__ NI_SOL int operator()(const Dop<1> &dop, int nr, double xold, double x, const double *y) {
    const double x_lo = fmin(xold, x) - xtol, x_hi = fmax(xold, x) + xtol;
    while (j < m) {
      const double xe = x_eval[j * step];
      if (xe < x_lo || xe > x_hi) break;
      y_eval[j * step] = dop.contd8(0, xe);
      j++;
    }
    return 1;
  }
};

// integrate_altitude_segment (clima_adiabat_altitude.f90:195-243) through solve_ivp_dop853 (clima_useful.f90:97-236):
// z = 0 at P_start, samples at the edges j_start, j_start + step, ..., j_end.  false = the integration failed.
__device__ NI_SOL bool altitude_segment(const AdiabatData &a, const AltitudeRhs &rhs, double P_start, const double *P_e2,
                                        double *z_e, int j_start, int j_end, int step) {
  Dop<1> dop;
  AltitudeSolout as;
  as.x_eval = P_e2 + j_start;
  as.y_eval = z_e + j_start;
  as.m = (j_end > j_start ? j_end - j_start : j_start - j_end) + 1;
  as.j = 0;
  as.step = step;
  const double big = fmax(1.0, fmax(P_start, fmax(P_e2[j_start], P_e2[j_end])));
  as.xtol = fmax(1.0e-12 * big, 10.0 * (nextafter(big, DBL_MAX) - big));
  while (as.j < as.m && fabs(as.x_eval[as.j * step] - P_start) <= as.xtol) {  // :155-159
    as.y_eval[as.j * step] = 0.0;
    as.j++;
  }
  if (as.j >= as.m) return true;
  double x = P_start, y[1] = {0.0};
  const int idid = dop_integrate<1>(dop, x, y, P_e2[j_end], a.rtol, a.atol, rhs, as);
  return idid >= 0 && as.j >= as.m;
}

// compute_altitude (clima_adiabat_altitude.f90:52-55, 97-193) on the prepared tables: the strictly-decreasing check
// of the second edge grid, then z at every edge -- from the surface upward (reference_pressure <= 0: the planet radius
// is at the surface), or from the reference pressure upward and downward and shifted back to zero at the surface
// (:107-169) -- and the extrapolated top edge.  Returns the column status (0, 8, 9 or 11).
__device__ NI_SOL int altitude_profile(const AdiabatData &a, int nz, double P_surf, const double *lp, const double *Tg,
                                       const double *mg, const double *P_e2, double *z_e) {
  double dmin = DBL_MAX;
  for (int k = 0; k < 2 * nz; k++) dmin = fmin(dmin, P_e2[k] - P_e2[k + 1]);
  const double ulp = nextafter(P_surf, DBL_MAX) - P_surf;  // spacing(P_e(1))
  if (dmin <= 10.0 * ulp) return 9;
  const AltitudeRhs rhs{lp, Tg, mg, nz, a.planet_mass, a.planet_radius};
  const int ne = 2 * nz + 1;
  const double Pr = a.reference_pressure;
  if (Pr <= 0.0) {
    z_e[0] = 0.0;
    if (!altitude_segment(a, rhs, P_surf, P_e2, z_e, 1, ne - 2, 1)) return 8;
  } else {
    if (Pr < a.P_top || Pr > P_surf) return 11;
    int k = 0;
    while (k < ne - 2 && !(Pr <= P_e2[k] && Pr >= P_e2[k + 1])) k++;
    int top, bot;
    if (Pr == P_e2[k]) {
      z_e[k] = 0.0;
      top = k + 1;
      bot = k - 1;
    } else if (Pr == P_e2[k + 1]) {
      z_e[k + 1] = 0.0;
      top = k + 2;
      bot = k;
    } else {
      top = k + 1;
      bot = k;
    }
    if (top <= ne - 2 && !altitude_segment(a, rhs, Pr, P_e2, z_e, top, ne - 2, 1)) return 8;
    if (bot >= 0 && !altitude_segment(a, rhs, Pr, P_e2, z_e, bot, 0, -1)) return 8;
    const double z_shift = z_e[0];  // keep the stored altitude surface-anchored (:164-166)
    for (int j = 0; j < ne - 1; j++) z_e[j] -= z_shift;
  }
  z_e[ne - 1] = z_e[ne - 2] + (z_e[ne - 2] - z_e[ne - 3]);
  return 0;
}

// AdiabatClimate_interpolate_particles (clima_adiabat.f90:1022-1044) + the particle part of
// copy_atm_to_radiative_grid (:745-750, :760-761) for the layers first, first + stride, ... of column c: densities and
// radii are 10 ** (linear interpolation in log10 P of the log10 values given to set_particle_density_and_radii,
// :1095-1121, whose table ends at log10(tiny) and log10(huge): every pressure is inside it).
__device__ void write_particles(const AdiabatData &a, const AdiabatOut &o, const int c, const double *P_m, const int first,
                                const int stride) {
  if (a.np <= 0 || !o.pd_r) return;
  const int nz = a.nz, n = a.npl;
  const bool dbl = a.double_radiative_grid != 0;
  const int nzr = dbl ? 2 * nz + 2 : nz;
  double *pd = o.pd_r + (size_t)nzr * a.np * c, *rd = o.rd_r + (size_t)nzr * a.np * c;
  for (int i = first; i < nz; i += stride) {
    const double x = nlog10(P_m[i]);
    int lo = 0, hi = n - 1;  // bracketing interval of the ascending grid (dintrv)
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (x >= a.part_x[mid]) lo = mid; else hi = mid;
    }
    const double q = (x - a.part_x[lo]) / (a.part_x[lo + 1] - a.part_x[lo]), p1 = 1.0 - q;
    const int reps = dbl ? (i == nz - 1 ? 4 : 2) : 1;
    const int j0 = dbl ? 2 * i : i;
    for (int k = 0; k < a.np; k++) {
      const double d = npow(10.0, p1 * a.part_d[lo + (size_t)n * k] + q * a.part_d[lo + 1 + (size_t)n * k]);
      const double r = npow(10.0, p1 * a.part_r[lo + (size_t)n * k] + q * a.part_r[lo + 1 + (size_t)n * k]);
      for (int t = 0; t < reps; t++) {
        pd[j0 + t + (size_t)nzr * k] = d;
        rd[j0 + t + (size_t)nzr * k] = r;
      }
    }
  }
}

// One column per thread.  Model-grid outputs are (nz, ncol) / (nz, ng, ncol) with nz fastest; the radiative
// grid (nzr = 2 nz + 2 when doubled, else nz) likewise.  `work` holds, per column, lp / Tg / mg (nz + 1 each),
// the second edge grid and its altitudes (2 nz + 1 each).
__global__ void __launch_bounds__(ADIABAT_THREADS)
make_profile_kernel(const AdiabatData *__restrict__ ap, int ncol, int lanes, const double *__restrict__ T_surf_in,
                    const double *__restrict__ P_i_surf, AdiabatOut o, double *__restrict__ work) {
  // `lanes` columns per warp (a power of two): a warp runs as long as its slowest column and a small batch
  // fills only a few of the 592 warp schedulers, so small batches spread their columns over more warps
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if ((t & 31) >= lanes) return;
  const int c = (t >> 5) * lanes + (t & 31);
  if (c >= ncol) return;
  const AdiabatData &a = *ap;
  const int ng = a.ng, nz = a.nz, ne = 2 * nz + 1;
  const double T_surf = T_surf_in[c];
  const double *Pin = P_i_surf + (size_t)ng * c;
  double *T_m = o.T + (size_t)nz * c, *P_m = o.P + (size_t)nz * c, *z_m = o.z + (size_t)nz * c,
         *dz_m = o.dz + (size_t)nz * c, *f_m = o.f_i + (size_t)nz * ng * c;
  double *N_surface = o.N_surface + (size_t)ng * c;
  int status = 0;
  Col col;
  col.a = ap;
  col.ng = ng;
  col.err = 0;
  col.nfcn = 0;
  o.P_trop[c] = -1.0;
  o.P_surf[c] = 0.0;

  // input checks (:124-139)
  const double T_trop = o.T_trop_in ? o.T_trop_in[c] : a.T_trop;
  col.T_trop = T_trop;
  if (T_surf < T_trop) status = 1;
  for (int i = 0; i < ng; i++)
    if (Pin[i] < 0.0 && !status) status = 2;
  if (T_trop < 0.0 && !status) status = 10;
  if (status) {
    o.status[c] = status;
    return;
  }

  // which gases condense at the surface (:199-224), the surface reservoir of those that do
  const double grav0 = gravity(a.planet_radius, a.planet_mass, 0.0);
  double f_cur[MAXG], P_surf = 0.0;
  for (int i = 0; i < ng; i++) {
    const double P_sat = a.has_sat[i] ? a.RH[i] * sat_pressure(a.sat[i], T_surf) : DBL_MAX;
    if (Pin[i] > P_sat) {
      f_cur[i] = P_sat;
      N_surface[i] = (Pin[i] - P_sat) / (a.mass[i] * grav0);
      col.cond[i] = 1;
    } else {
      f_cur[i] = Pin[i];
      N_surface[i] = 0.0;
      col.cond[i] = 0;
    }
    P_surf += f_cur[i];
  }
  o.P_surf[c] = P_surf;
  if (a.P_top > P_surf) {
    o.status[c] = 3;
    return;
  }
  for (int i = 0; i < ng; i++) f_cur[i] /= P_surf;
  update_f_i_dry(col, P_surf, f_cur);
  for (int i = 0; i < ng; i++) o.f_i_surf[(size_t)ng * c + i] = f_cur[i];

  ProfileSolout so;
  so.col = &col;
  so.ne = ne;
  so.j = 1;
  so.l0 = nlog10(P_surf);
  so.range = nlog10(a.P_top) - so.l0;
  so.P_surf = P_surf;
  so.Pj = so.P_edge(1);
  so.reason = 0;
  so.z_negative = false;
  so.T_m = T_m;
  so.f_m = f_m;
  so.nz = nz;
  for (int i = 0; i < nz; i++) P_m[i] = so.P_edge(2 * i + 1);

  // integrate (:274-353): restart after every "gas reached saturation" event with that gas condensing
  {
    Dop<2> dop;
    ProfileRhs rhs{&col};
    double u[2] = {T_surf, 0.0}, Pn = P_surf;
    for (;;) {
      so.reason = 0;
      const int idid = dop_integrate<2>(dop, Pn, u, a.P_top, a.rtol, a.atol, rhs, so);
      if (col.err) {
        status = col.err;
        break;
      }
      if (idid < 0) {
        status = 4;
        break;
      }
      if (so.reason != 2) break;
      Pn = so.P_root;
      u[0] = so.u_root[0];
      u[1] = so.u_root[1];
      mixing_ratios(col, Pn, u[0], f_cur);
      col.cond[so.ind_sat] = 1;
      update_f_i_dry(col, Pn, f_cur);
    }
  }
  if (!status && so.reason == 1) {  // isothermal stratosphere above the tropopause (:331-351)
    mixing_ratios(col, so.P_root, so.u_root[0], f_cur);
    o.P_trop[c] = so.P_root;
    double mubar = 0.0;
    for (int i = 0; i < ng; i++) mubar += f_cur[i] * a.mass[i];
    for (int k = so.j; k < ne; k++)
      so.sample(k, T_trop,
                altitude_vs_P(so.P_edge(k), T_trop, mubar, so.P_root, so.u_root[1], a.planet_mass, a.planet_radius),
                f_cur);
  }
  if (!status && so.z_negative) status = 7;
  o.nfcn[c] = col.nfcn;
  if (status) {
    o.status[c] = status;
    return;
  }

  // compute_altitude (clima_adiabat_altitude.f90:41-193)
  double *lp = work + (size_t)c * (3 * (nz + 1) + 2 * ne), *Tg = lp + nz + 1, *mg = Tg + nz + 1, *P_e2 = mg + nz + 1,
         *z_e = P_e2 + ne;
  lp[0] = nlog10(P_surf);
  Tg[0] = T_surf;
  {
    double m0 = 0.0;
    for (int g = 0; g < ng; g++) m0 += o.f_i_surf[(size_t)ng * c + g] * a.mass[g];
    mg[0] = m0;
  }
  P_e2[0] = P_surf;
  for (int i = 0; i < nz; i++) {
    lp[i + 1] = nlog10(P_m[i]);
    Tg[i + 1] = T_m[i];
    double m = 0.0;
    for (int g = 0; g < ng; g++) m += f_m[i + (size_t)nz * g] * a.mass[g];
    mg[i + 1] = m;
    P_e2[2 * i + 1] = P_m[i];
    if (i < nz - 1) P_e2[2 * i + 2] = sqrt(P_m[i] * P_m[i + 1]);
  }
  P_e2[2 * nz] = a.P_top;
  status = altitude_profile(a, nz, P_surf, lp, Tg, mg, P_e2, z_e);
  o.status[c] = status;
  if (status) return;
  for (int i = 0; i < nz; i++) {
    z_m[i] = z_e[2 * i + 1];
    dz_m[i] = z_e[2 * i + 2] - z_e[2 * i];
  }

  // densities (clima_adiabat.f90:446-449) and copy_atm_to_radiative_grid (:729-773), pressures in bars
  if (o.T_r) {
    const int nzr = a.double_radiative_grid ? 2 * nz + 2 : nz;
    double *T_r = o.T_r + (size_t)nzr * c, *P_r = o.P_r + (size_t)nzr * c, *dz_r = o.dz_r + (size_t)nzr * c,
           *dens_r = o.dens_r + (size_t)nzr * ng * c;
    for (int i = 0; i < nz; i++) {
      const double density = P_m[i] / (K_BOLTZ * T_m[i]);
      const int reps = a.double_radiative_grid ? (i == nz - 1 ? 4 : 2) : 1;
      const int j0 = a.double_radiative_grid ? 2 * i : i;
      const double dzr = a.double_radiative_grid ? 0.5 * dz_m[i] : dz_m[i];
      for (int r = 0; r < reps; r++) {
        T_r[j0 + r] = T_m[i];
        P_r[j0 + r] = P_m[i] / 1.0e6;
        dz_r[j0 + r] = dzr;
        for (int g = 0; g < ng; g++) dens_r[j0 + r + (size_t)nzr * g] = f_m[i + (size_t)nz * g] * density;
      }
    }
    o.Tsurf_r[c] = T_surf;
    write_particles(a, o, c, P_m, 0, 1);
  }
}

#include "adiabat_group.cuh"

}  // namespace adiabat

size_t rt_adiabat_work_doubles(int nz, int ncol) { return (size_t)ncol * (3 * (nz + 1) + 2 * (2 * nz + 1)); }

void rt_launch_make_profile(const AdiabatData *a, int ng, int ncol, int num_sms, const double *T_surf,
                            const double *P_i_surf, const AdiabatOut &o, double *work, cudaStream_t st) {
  // Both kernels are bound by the latency of one column's chain of steps, and columns of a warp sit at different
  // places of the integrator (ncu: 2.3 of 4 active), so a small batch spreads over more warps: a warp takes more
  // columns only as the batch grows beyond ADIABAT_WARPS_PER_SM warps per SM.
  const char *e = getenv("CLIMA_RT_ADIABAT_WARPS_PER_SM");
  const int wps = e && atoi(e) > 0 ? atoi(e) : ADIABAT_WARPS_PER_SM;
  // Which kernel: one lane per gas shortens a column's chain (4096 columns: 17.8 against 27.1 ms; 256: 10.2 against
  // 15.8) but occupies 8 lanes per column, so beyond about 64 columns per SM the one-thread-per-column kernel
  // wins on throughput (16384 columns: 31.6 against 42.9 ms; the crossover was measured near 10 000 columns).
  // CLIMA_RT_ADIABAT=thread|group forces one (read per call: the tests run both in one process).
  const char *m = getenv("CLIMA_RT_ADIABAT");
  const bool per_thread = m && (!strcmp(m, "thread") || !strcmp(m, "group")) ? !strcmp(m, "thread")
                                                                             : ncol > ADIABAT_GROUP_MAX_COLS_PER_SM * num_sms;
  const int wpb = ADIABAT_THREADS / 32;
  if (per_thread) {  // one thread per column
    int lanes = 32;
    while (lanes > 1 && (size_t)ncol < (size_t)wps * num_sms * lanes) lanes >>= 1;
    const int warps = (ncol + lanes - 1) / lanes;
    adiabat::make_profile_kernel<<<(warps + wpb - 1) / wpb, ADIABAT_THREADS, 0, st>>>(a, ncol, lanes, T_surf, P_i_surf,
                                                                                    o, work);
    return;
  }
  // one lane per gas: 8 lanes per column up to 8 gases (4 columns per warp at most), 16 beyond
  const int gs = ng <= 8 ? 8 : 16;
  int gpw = 32 / gs;
  while (gpw > 1 && (size_t)ncol < (size_t)wps * num_sms * gpw) gpw >>= 1;
  const int warps = (ncol + gpw - 1) / gpw;
  const int blocks = (warps + wpb - 1) / wpb;
  if (gs == 8)
    adiabat::make_profile_group_kernel<8><<<blocks, ADIABAT_THREADS, 0, st>>>(a, ncol, gpw, T_surf, P_i_surf, o, work);
  else
    adiabat::make_profile_group_kernel<16><<<blocks, ADIABAT_THREADS, 0, st>>>(a, ncol, gpw, T_surf, P_i_surf, o, work);
}

// ---------------------------------------------------------------------------------------------------------
// make_profile with its per-column work arrays in a STATIC device array instead of a cudaMalloc'ed buffer
// (clima_rt_make_profile_batch_noalloc: for callers that may not allocate -- CUDA graph capture, serving loops).
// The array is held by every GPU that loads the library until the process exits; every call shares it, so calls
// that overlap on the GPU race.  Nothing in it has to start at zero: a column writes lp / Tg / mg / P_e2 / z_e
// before it reads them.
// ---------------------------------------------------------------------------------------------------------
static __device__ __align__(256) double g_noalloc_work[RT_NOALLOC_MAX_WORK_DOUBLES];

size_t rt_noalloc_max_work_doubles() { return (size_t)RT_NOALLOC_MAX_WORK_DOUBLES; }

// In host code `g_noalloc_work` is a host-side stand-in, not a device address: the address the kernel gets comes
// from cudaGetSymbolAddress, once per device, with that device current.  Returns nullptr and sets *why on failure.
static double *noalloc_work_ptr(const char **why) {
  static double *ptr[64] = {};
  static std::mutex mu;
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) {
    *why = "make_profile (no-alloc): cudaGetDevice failed";
    return nullptr;
  }
  std::lock_guard<std::mutex> lock(mu);
  if (!ptr[dev]) {
    double *p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress((void **)&p, g_noalloc_work);
    if (e != cudaSuccess || !p) {
      *why = "make_profile (no-alloc): cudaGetSymbolAddress of the static scratch array failed";
      return nullptr;
    }
    ptr[dev] = p;
  }
  return ptr[dev];
}

// 0 = launched; otherwise *why says what is wrong (nothing was launched).  The current device must be the one the
// pointers live on.
int rt_launch_make_profile_noalloc(const AdiabatData *a, int ng, int nz, int ncol, int num_sms, const double *T_surf,
                                   const double *P_i_surf, const AdiabatOut &o, cudaStream_t st, const char **why) {
  if (rt_adiabat_work_doubles(nz, ncol) > rt_noalloc_max_work_doubles()) {
    *why = "make_profile (no-alloc): ncol * (7 nz + 5) exceeds the static scratch array";
    return 1;
  }
  double *work = noalloc_work_ptr(why);
  if (!work) return 2;
  rt_launch_make_profile(a, ng, ncol, num_sms, T_surf, P_i_surf, o, work, st);
  return 0;
}
