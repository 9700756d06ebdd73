// prep.cu -- per-column prologue of OpticalProperties_compute_opacity
// (src/radtran/clima_radtran_types.f90:600-633) plus everything that is independent of the
// wavelength bin and can therefore be hoisted out of the per-bin hot loop:
//   * unique-layer slots: layer j shares the slot of layer j-1 when every input of the two layers
//     is bitwise equal (the doubled grid + ghost layers of src/adiabat/clima_adiabat.f90:729-773
//     are exact copies).  This generalises `pair_reuse` (:621-632, even j only, is_close 1e-12):
//     identical inputs give identical outputs, so sharing is exact.
//   * foreign_col (:608-618)
//   * bracket index + fraction of every 1-D/2-D table interpolation (dintrv + q of
//     linear_interpolation_module.F90:233-271,297-337): these depend on (P,T,radius) only, the
//     reference redoes the search for each of the nw bins.
// One block per column, one thread per layer.
#include "rt_device.cuh"
#include "rt_kernels.h"

namespace {

__device__ __forceinline__ bool same_bits(double a, double b) {
  return __double_as_longlong(a) == __double_as_longlong(b);
}

__global__ void prep_kernel(const DevTables tb, const DevChunk ch) {
  extern __shared__ int sm[];  // [nz] dup flags -> slot ids
  const int c = blockIdx.x;
  const int nz = ch.nz, ng = tb.ng, np = tb.np;
  const double *T = ch.T + (size_t)c * nz;
  const double *P = ch.P + (size_t)c * nz;
  const double *dz = ch.dz + (size_t)c * nz;
  const double *dens = ch.dens + (size_t)c * nz * ng;
  const double *pd = ch.pdens ? ch.pdens + (size_t)c * nz * np : nullptr;
  const double *rd = ch.radii ? ch.radii + (size_t)c * nz * np : nullptr;

  for (int j = threadIdx.x; j < nz; j += blockDim.x) {
    bool dup = false;
    if (j > 0) {
      dup = same_bits(T[j], T[j - 1]) && same_bits(P[j], P[j - 1]) && same_bits(dz[j], dz[j - 1]);
      for (int i = 0; i < ng && dup; i++)
        dup = same_bits(dens[j + (size_t)i * nz], dens[j - 1 + (size_t)i * nz]);
      if (rd)
        for (int i = 0; i < np && dup; i++)
          dup = same_bits(rd[j + (size_t)i * nz], rd[j - 1 + (size_t)i * nz]) &&
                same_bits(pd[j + (size_t)i * nz], pd[j - 1 + (size_t)i * nz]);
    }
    sm[j] = dup ? 0 : 1;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    int s = -1;
    for (int j = 0; j < nz; j++) {
      if (sm[j]) {
        s++;
        ch.layer_of_slot[(size_t)c * nz + s] = j;
      }
      sm[j] = sm[j] ? s : -1 - s;  // representative: slot ; duplicate: -(1+slot)
    }
    ch.nslots[c] = s + 1;
  }
  __syncthreads();

  const int nbr = rt_nbr(tb);
  for (int j = threadIdx.x; j < nz; j += blockDim.x) {
    const int v = sm[j];
    const int slot = v >= 0 ? v : -1 - v;
    ch.slot_of_layer[(size_t)c * nz + j] = slot;
    if (v < 0) continue;  // duplicate layer: nothing more to do
    int *ib = ch.ibr + ((size_t)c * nz + slot) * nbr;
    double *qb = ch.qbr + ((size_t)c * nz + slot) * nbr;
    const double Tj = T[j];
    const double log10P = log10(P[j]);  // :604
    int r = 0;
    // k-distributions: clamp then bracket (:654-657)
    for (int i = 0; i < tb.nk; i++) {
      const DevK &k = tb.k[i];
      double PP = fmin(fmax(log10P, k.logP[0]), k.logP[k.np - 1]);
      double TT = fmin(fmax(Tj, k.T[0]), k.T[k.nt - 1]);
      int ix = rt_bracket(k.logP, k.np, PP);
      int iy = rt_bracket(k.T, k.nt, TT);
      ib[r] = ix;
      qb[r++] = (PP - k.logP[ix]) / (k.logP[ix + 1] - k.logP[ix]);
      ib[r] = iy;
      qb[r++] = (TT - k.T[iy]) / (k.T[iy + 1] - k.T[iy]);
    }
    // CIA / photolysis (interpolate_Xsection :890-917)
    for (int i = 0; i < tb.ncia + tb.npxs; i++) {
      const DevXS &x = i < tb.ncia ? tb.cia[i] : tb.pxs[i - tb.ncia];
      if (x.dim == 1) {
        double TT = fmin(fmax(Tj, x.temp[0]), x.temp[x.nt - 1]);
        int it = rt_bracket(x.temp, x.nt, TT);
        ib[r] = it;
        qb[r++] = (TT - x.temp[it]) / (x.temp[it + 1] - x.temp[it]);
      } else {
        ib[r] = 0;
        qb[r++] = 0.0;
      }
    }
    // water continuum (:919-945)
    if (tb.has_cont) {
      double TT = fmin(fmax(Tj, tb.cont_temp[0]), tb.cont_temp[tb.cont_nt - 1]);
      int it = rt_bracket(tb.cont_temp, tb.cont_nt, TT);
      ib[r] = it;
      qb[r++] = (TT - tb.cont_temp[it]) / (tb.cont_temp[it + 1] - tb.cont_temp[it]);
      double fc = 0.0;  // :608-618
      for (int i = 0; i < ng; i++)
        if (i != tb.cont_LH2O - 1) fc = fc + dens[j + (size_t)i * nz] * dz[j];
      ch.foreign_col[(size_t)c * nz + slot] = fc;
    } else {
      ib[r] = 0;
      qb[r++] = 0.0;
    }
    // particles (:947-983): radius clamped, out-of-range raises the status bit
    for (int i = 0; i < tb.npart; i++) {
      const DevPart &p = tb.part[i];
      double rp = rd[j + (size_t)(p.p_ind - 1) * nz];
      const double r_min = p.radii[0], r_max = p.radii[p.nr - 1];
      if (rp < r_min || rp > r_max) {
        rp = fmin(fmax(rp, r_min), r_max);
        atomicOr(ch.status + c, 1);
      }
      int ir = rt_bracket(p.radii, p.nr, rp);
      ib[r] = ir;
      qb[r++] = (rp - p.radii[ir]) / (p.radii[ir + 1] - p.radii[ir]);
    }
    // custom optical properties (:546-556): interpolated (extrapolated) in log10 P[dyn/cm2]
    if (tb.has_custom) {
      double lp = log10(P[j] * 1.0e6);  // :605
      int ip = rt_bracket(tb.c_log10P, tb.custom_nP, lp);
      ib[r] = ip;
      qb[r++] = (lp - tb.c_log10P[ip]) / (tb.c_log10P[ip + 1] - tb.c_log10P[ip]);
    } else {
      ib[r] = 0;
      qb[r++] = 0.0;
    }
  }
}

// Are the 4 adjacent columns that share a RORR warp alike?  Thread per group of 4 columns: the
// k-gas column amounts of the ground layer and of the middle layer must agree within 10 % over the group
// (dissim[0]: unit order of the sorting kernels) and within 25 % between NEIGHBOURS, the first column of the
// next group included (dissim[1]: is the staircase kernel worth a try; what it then reports decides, capi.cu).
__global__ void similarity_kernel(const DevTables tb, const DevChunk ch) {
  const int g4 = blockIdx.x * blockDim.x + threadIdx.x;
  const int c0 = g4 * 4;
  if (c0 >= ch.ncol) return;
  const int nz = ch.nz, ng = tb.ng;
  bool alike = true, near = true;
  for (int jj = 0; jj < tb.nk; jj++) {
    const int sp = tb.k[jj].sp - 1;
    for (int kk = 0; kk < 2; kk++) {
      const int k = kk ? nz / 2 : 0;
      double lo = 0.0, hi = 0.0, prev = 0.0;
      for (int c = c0; c < c0 + 5 && c < ch.ncol; c++) {
        const double v = ch.dens[(size_t)c * nz * ng + k + (size_t)sp * nz] * ch.dz[(size_t)c * nz + k];
        if (c > c0 && !(fmax(v, prev) <= 1.25 * fmin(v, prev))) near = false;
        prev = v;
        if (c == c0 + 4) break;
        if (c == c0) { lo = hi = v; } else { lo = fmin(lo, v); hi = fmax(hi, v); }
      }
      if (!(hi <= 1.1 * lo)) alike = false;
    }
  }
  if (!alike) atomicAdd(ch.dissim, 1);
  if (!near) atomicAdd(ch.dissim + 1, 1);
}

}  // namespace

void rt_launch_similarity(const DevTables &tb, const DevChunk &ch, cudaStream_t st) {
  cudaMemsetAsync(ch.dissim, 0, 2 * sizeof(int), st);
  const int groups = (ch.ncol + 3) / 4;
  similarity_kernel<<<(groups + 127) / 128, 128, 0, st>>>(tb, ch);
}

void rt_launch_prep(const DevTables &tb, const DevChunk &ch, cudaStream_t st) {
  int threads = 128;
  prep_kernel<<<ch.ncol, threads, ch.nz * sizeof(int), st>>>(tb, ch);
}
