// continua_impl.cuh -- Rayleigh, CIA, photolysis, water continuum, particles, custom opacity and the
// asymmetry parameter of one (column, bin, unique layer slot): everything of
// OpticalProperties_compute_opacity (src/radtran/clima_radtran_types.f90:664-757) that does not
// depend on the g-point.  Runs as its own small kernel (optics_rorr.cu) so that the RORR kernel
// carries neither this code (instruction-cache footprint) nor its results (registers).
#pragma once
#include "rt_device.cuh"

__device__ __forceinline__ double shfl_xor_d(double v, int lanemask) {
  return __shfl_xor_sync(0xffffffffu, v, lanemask);
}

// ---------------------------------------------------------------------------------------------
struct Continua {
  double tausg, taua, taup, tausp, tauc, tausc, gt;
};

template <int LPU>
__device__ __forceinline__ double unit_sum(double v) {
#pragma unroll
  for (int m = LPU / 2; m >= 1; m >>= 1) v = __dadd_rn(v, shfl_xor_d(v, m));
  return v;
}

// Rayleigh, CIA, photolysis, water continuum, particles, custom opacity and the asymmetry
// parameter of one (column c, bin l, slot) -- A.3, clima_radtran_types.f90:664-757.
// Items (nray Rayleigh, ncia CIA, npxs cross sections, 2 continuum terms) are dealt to the unit's
// lanes round-robin; every lane ends with the same totals.
template <int LPU>
__device__ __forceinline__ Continua rt_continua(const DevTables &tb, const DevChunk &ch, const int c,
                                                const int l, const int slot, const int k,
                                                const int ul) {
  const int nw = tb.nw, nz = ch.nz, ng = tb.ng;
  const double *dens = ch.dens + (size_t)c * nz * ng + k;  // dens[i*nz]
  const double dzk = ch.dz[(size_t)c * nz + k];
  const int nbr = rt_nbr(tb);
  const int *ib = ch.ibr + ((size_t)c * nz + slot) * nbr;
  const double *qb = ch.qbr + ((size_t)c * nz + slot) * nbr;
  const int rb_xs = 2 * tb.nk;
  const int rb_cont = rb_xs + tb.ncia + tb.npxs;
  const int rb_part = rb_cont + 1;
  const int rb_cust = rb_part + tb.npart;
  Continua o;
  double tausg = 0.0, taua = 0.0;
  const int n_xs = tb.ncia + tb.npxs;
  const int nitems = tb.nray + n_xs + (tb.has_cont ? 2 : 0);
#pragma unroll 1
  for (int it = ul; it < nitems; it += LPU) {
    if (it < tb.nray) {  // Rayleigh (:686-693)
      const double col = dens[(size_t)(tb.ray_sp[it] - 1) * nz] * dzk;
      tausg = tausg + tb.ray_xs[l + (size_t)it * nw] * col;
    } else if (it < tb.nray + n_xs) {  // CIA (:696-704), photolysis (:707-713)
      const int i = it - tb.nray;
      const bool is_cia = i < tb.ncia;
      const DevXS &x = is_cia ? tb.cia[i] : tb.pxs[i - tb.ncia];
      double xs;
      if (x.dim == 0) {
        xs = x.data[l];
      } else {
        const int t0 = ib[rb_xs + i];
        const double q1 = qb[rb_xs + i], p1 = 1.0 - q1;
        const double *f = x.data + (size_t)l * x.nt;
        xs = rt_ten2power(p1 * f[t0] + q1 * f[t0 + 1]);
      }
      if (is_cia)
        taua = taua + xs * dens[(size_t)(x.sp1 - 1) * nz] * dens[(size_t)(x.sp2 - 1) * nz] * dzk;
      else
        taua = taua + xs * (dens[(size_t)(x.sp1 - 1) * nz] * dzk);
    } else {  // water continuum (:716-723): self term, then foreign term
      const bool self = (it == tb.nray + n_xs);
      const int t0 = ib[rb_cont];
      const double q1 = qb[rb_cont], p1 = 1.0 - q1;
      const double *f = (self ? tb.cont_H2O : tb.cont_foreign) + (size_t)l * tb.cont_nt;
      const double xs = rt_ten2power(p1 * f[t0] + q1 * f[t0 + 1]);
      const double dw = dens[(size_t)(tb.cont_LH2O - 1) * nz];
      if (self) taua = taua + xs * dw * (dw * dzk);
      else taua = taua + xs * dw * ch.foreign_col[(size_t)c * nz + slot];
    }
  }
  tausg = unit_sum<LPU>(tausg);
  taua = unit_sum<LPU>(taua);
  // particles (:733-744)
  double tausp = 0.0, taup = 0.0;
  double tausp_1[RT_MAXPART], gtp[RT_MAXPART];
  if (tb.npart > 0) {
    const double *rad = ch.radii + (size_t)c * nz * tb.np + k;
    const double *pdn = ch.pdens + (size_t)c * nz * tb.np + k;
#pragma unroll 1
    for (int i = 0; i < tb.npart; i++) {
      const DevPart &p = tb.part[i];
      const int ir = ib[rb_part + i];
      const double q1 = qb[rb_part + i], p1 = 1.0 - q1;
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
  // custom optical properties (:726-730, 546-572)
  double tauc = RT_TINY, w0c = RT_TINY, g0c = RT_TINY;
  if (tb.has_custom) {
    const int ip = ib[rb_cust];
    const double q1 = qb[rb_cust], p1 = 1.0 - q1;
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

