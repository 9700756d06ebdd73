// rt_device.cuh -- device-side data model shared by all kernels of the RT hot path.
//
// HBM layout (all FP64, per chunk of `ncol` columns; "slot" = index of a UNIQUE layer of a column,
// ground-first; duplicated layers of the doubled radiative grid (src/adiabat/clima_adiabat.f90:
// 729-773) share one slot, generalising `pair_reuse` of clima_radtran_types.f90:621-632):
//
//   tau  [col][bin l][slot][g]      g fastest: one 8*G-byte line per (col,l,slot); the two-stream
//   w0   [col][bin l][slot][g]      threads (one per g) read it fully coalesced
//   gt   [col][bin l][slot]
//   tband[col][bin l][slot]
//   bpl  [col][ir bin ll][level n]  Planck, top-first, level nz = surface
//   fup_a/fdn_a/amean [col][ll][level i]  ground-first == Fortran (nz+1,nw_ch,ncol) column-major
//
// Tables are read-only and L2-resident (tens of MB); they keep the reference's own layout
// log10k(G,np,nt,nw) (clima_radtran_types_create.f90:1291-1358) because G-fastest is exactly what
// a lane-per-g interpolation wants.
#pragma once
#include <cstdint>

#define RT_MAXK 12
#define RT_MAXCIA 32
#define RT_MAXRAY 16
#define RT_MAXPXS 16
#define RT_MAXPART 8
#define RT_MAXZEN 16

// src/clima_const.f90:9-21
#define RT_K_BOLTZ_SI 1.380649e-23
#define RT_PLANK 6.62607004e-34
#define RT_C_LIGHT 299792458.0
#define RT_PI 3.14159265358979323846e0
// src/radtran/clima_radtran_types.f90:9-11
#define RT_MAX_W0 0.99999
#define RT_MAX_GT 0.999999
#define RT_TAU_MIN 1.0e-20
#define RT_LN10 2.302585092994045684017991454684364  // log(10d0), clima_eqns.f90:78
#define RT_TINY 2.2250738585072014e-308              // tiny(0d0), clima_radtran_types.f90:558-563

struct DevK {  // type Ktable
  const double *logP, *T, *log10k;
  int sp, np, nt, pad;
};
struct DevXS {  // type Xsection (CIA: sp1,sp2; photolysis: sp1)
  const double *temp, *data;
  int sp1, sp2, dim, nt;
};
struct DevPart {  // type ParticleXsection
  const double *radii, *w0, *qext, *gt;
  int p_ind, nr;
};

struct DevTables {
  int nw, ng, np, nk, G, ncia, nray, npxs, npart;
  int has_cont, cont_LH2O, cont_nt;
  int ir_start, ir_end, sol_start, sol_end;  // 0-based inclusive bin ranges
  int has_custom, custom_nP, rows_sorted, pad0;
  DevK k[RT_MAXK];
  DevXS cia[RT_MAXCIA];
  DevXS pxs[RT_MAXPXS];
  DevPart part[RT_MAXPART];
  int ray_sp[RT_MAXRAY];
  const double *ray_xs;  // (nw,nray)
  const double *cont_temp, *cont_H2O, *cont_foreign;
  const double *wavl, *freq;             // (nw+1)
  const double *wbin, *wbin_e, *inv_dw;  // (G),(G+1),(G): Ksettings + 1/(wbin_e(g+1)-wbin_e(g))
  const double *c_log10P, *c_dtau, *c_w0, *c_g0;  // custom optical properties (nP),(nP,nw)
  const double *wxy;  // (P*P*4): wbin(i)*wbin(j) at [(i*P+j)*4 + copy], P = padded ngauss, 0 for padding
  double wq[32];      // wbin by value (zero padded): constant-bank operands of the staircase kernel
};

// number of bracket records per (col,slot): [k gases: 2 each][cia][pxs][cont][part][custom]
__host__ __device__ inline int rt_nbr(const DevTables &tb) {
  return 2 * tb.nk + tb.ncia + tb.npxs + 1 + tb.npart + 1;
}

struct DevParams {  // public members of type Radtran (clima_radtran.f90:51-68)
  double diurnal_fac, photon_scale_factor, ir_tau_min;
  int has_hard_surface, nzen;
  double zen_u[RT_MAXZEN], zen_w[RT_MAXZEN];
  const double *albedo, *emiss, *photons_sol;  // device (nw_sol),(nw_ir),(nw_sol)
};

// One chunk of columns resident on the device.
struct DevChunk {
  int ncol, nz, nslot_cap;
  // inputs, ground-first, column index slowest (exactly the C-ABI layout)
  const double *T_surface, *T, *P, *dens, *dz, *pdens, *radii;
  // prep results
  int *slot_of_layer;  // [col][nz]
  int *layer_of_slot;  // [col][nz]
  int *nslots;         // [col]
  int *ibr;            // [col][slot][nbr]   bracket indices (0-based left index)
  double *qbr;         // [col][slot][nbr]   bracket fractions q
  double *foreign_col; // [col][slot]
  // opacity (OpticalPropertiesResult, clima_radtran_types.f90:242-247), slot-indexed
  double *tau, *w0, *gt, *tband;
  double *cont;  // [3][col][bin l][slot]: g-independent terms handed from continua_kernel to the RORR epilogue
  int *status;  // [col]
  // [0]: groups of 4 adjacent columns whose k-gas amounts differ (counted by similarity_kernel);
  // the RORR kernels choose their unit order from it
  int *dissim;
  // staircase RORR kernel bookkeeping (accumulates over calls): [0] warp-merges, [1] warp-passes,
  // [2] restarts from bisection, [3] units handed to the exact kernel, [4] cold starts (warp-merges)
  unsigned long long *cut_stats;
};

// ---- small device helpers shared by kernels -------------------------------------------------
// dintrv bracket semantics, linear_interpolation_module.F90:348-350 (0-based left index)
__device__ __forceinline__ int rt_bracket(const double *__restrict__ xt, int n, double x) {
  if (x >= xt[n - 1]) return n - 2;
  if (x < xt[0]) return 0;
  int lo = 0, hi = n - 1;
  while (hi - lo > 1) {
    int mid = (lo + hi) >> 1;
    if (x < xt[mid]) hi = mid; else lo = mid;
  }
  return lo;
}
// ten2power, src/clima_eqns.f90:75-80
__device__ __forceinline__ double rt_ten2power(double y) { return exp(y * RT_LN10); }
// exp(x) for |x| < 700 without the special-case handling of the library routine (no overflow,
// underflow, NaN paths): k = rint(x*log2e) by the 1.5*2^52 trick, Cody-Waite reduction with a
// two-part ln2, degree-13 Taylor polynomial on |r| <= ln2/2 (remainder 4e-18), scaling by an
// exponent-field add.  <= 1 ulp (checked against libm on 2e6 arguments); 20 instructions instead
// of ~49 -- the RORR kernel is instruction-issue bound and evaluates 2 of these per lane and gas.
__device__ __forceinline__ double rt_exp_lean(const double x) {
  const double t = __fma_rn(x, 1.4426950408889634074, 6755399441055744.0);
  const int k = __double2loint(t);
  const double kf = __dsub_rn(t, 6755399441055744.0);
  double r = __fma_rn(kf, -6.93147180369123816490e-01, x);
  r = __fma_rn(kf, -1.90821492927058770002e-10, r);
  double p = 1.6059043836821613e-10;        // 1/13!
  p = __fma_rn(p, r, 2.08767569878681e-09);   // 1/12!
  p = __fma_rn(p, r, 2.505210838544172e-08);  // 1/11!
  p = __fma_rn(p, r, 2.755731922398589e-07);  // 1/10!
  p = __fma_rn(p, r, 2.7557319223985893e-06); // 1/9!
  p = __fma_rn(p, r, 2.48015873015873e-05);   // 1/8!
  p = __fma_rn(p, r, 1.984126984126984e-04);  // 1/7!
  p = __fma_rn(p, r, 1.388888888888889e-03);  // 1/6!
  p = __fma_rn(p, r, 8.333333333333333e-03);  // 1/5!
  p = __fma_rn(p, r, 4.1666666666666664e-02); // 1/4!
  p = __fma_rn(p, r, 1.6666666666666666e-01); // 1/3!
  p = __fma_rn(p, r, 0.5);
  p = __fma_rn(p, r, 1.0);
  p = __fma_rn(p, r, 1.0);
  return __hiloint2double(__double2hiint(p) + (k << 20), __double2loint(p));
}
// 1/d for d well inside the normal range (Thomas pivots, e1 in [1,2], gamma sums): the hardware
// seed (2^-23) and two Newton steps, ~1 ulp, no denormal/overflow slow path.  The two-stream kernels
// are FP64-pipe / issue limited: the IEEE division costs ~20 instructions, this 5.
__device__ __forceinline__ double rt_rcp_fast(const double d) {
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(d));
  double e = __fma_rn(-d, r, 1.0);
  r = __fma_rn(r, e, r);
  e = __fma_rn(-d, r, 1.0);
  return __fma_rn(r, e, r);
}
// exp(x) for x <= 0 (attenuation factors); below -700 the result is held at exp(-700) ~ 1e-304,
// which no sum of the solver can tell from 0
__device__ __forceinline__ double rt_exp_neg(const double x) { return rt_exp_lean(x < -700.0 ? -700.0 : x); }
// ten2power for arguments known to be far from the overflow/underflow limits (k-table values:
// log10 k in [-60, 10])
__device__ __forceinline__ double rt_ten2power_lean(double y) { return rt_exp_lean(y * RT_LN10); }
// out-of-range arguments (tables holding -inf or absurd values): the library routine, kept out of
// line so that the hot code does not carry it
static __device__ __noinline__ double rt_ten2power_cold(double y) { return exp(y * RT_LN10); }
// planck_fcn, src/clima_eqns.f90:64-73
__device__ __forceinline__ double rt_planck(double nu, double T) {
  return 1.0e3 * ((2.0 * RT_PLANK * (nu * nu * nu)) / (RT_C_LIGHT * RT_C_LIGHT)) *
         ((1.0) / (exp((RT_PLANK * nu) / (RT_K_BOLTZ_SI * T)) - 1.0));
}
