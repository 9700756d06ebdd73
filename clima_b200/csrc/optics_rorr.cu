// optics_rorr.cu -- launch sequence of the optics + RORR stage and the small layout kernels around
// it.  The kernel itself lives in rorr_impl.cuh (instantiated by optics_rorr_fast.cu /
// optics_rorr_exact.cu).
#include <algorithm>

#include "continua_impl.cuh"
#include "rt_device.cuh"
#include "rt_kernels.h"

void rorr_launch_fast(const DevTables &tb, const DevChunk &ch, int *worklist, int *wcount, int wcap,
                      long long units, cudaStream_t st);
void rorr_launch_exact(const DevTables &tb, const DevChunk &ch, int *worklist, int *wcount, int wcap,
                       long long units, cudaStream_t st);
void rorr_launch_cut(const DevTables &tb, const DevChunk &ch, int mode, int Lc, int *worklist, int *wcount,
                     int wcap, cudaStream_t st);

namespace {

// g-independent opacity terms, one thread per (column, bin, unique layer slot), slot fastest.
// Writes the asymmetry parameter and the three numbers the RORR epilogue (:856-886) needs:
//   cont[0*U + u] = (tausg + taua) + taup     (the leading terms of the total tau, in that order)
//   cont[1*U + u] = tauc
//   cont[2*U + u] = tausg + tausp + tausc     (numerator of w0)
__global__ void __launch_bounds__(256) continua_kernel(const __grid_constant__ DevTables tb,
                                                       const __grid_constant__ DevChunk ch) {
  const int cap = ch.nslot_cap, nw = tb.nw, nz = ch.nz;
  const long long U = (long long)ch.ncol * nw * cap;
  for (long long u = blockIdx.x * (long long)blockDim.x + threadIdx.x; u < U;
       u += (long long)gridDim.x * blockDim.x) {
    const int slot = (int)(u % cap);
    const int l = (int)((u / cap) % nw);
    const int c = (int)(u / ((long long)cap * nw));
    if (slot >= ch.nslots[c]) continue;
    const int k = ch.layer_of_slot[(size_t)c * nz + slot];
    const Continua q = rt_continua<1>(tb, ch, c, l, slot, k, 0);
    ch.cont[u] = __dadd_rn(__dadd_rn(q.tausg, q.taua), q.taup);
    ch.cont[U + u] = q.tauc;
    ch.cont[2 * U + u] = q.tausg + q.tausp + q.tausc;
    ch.gt[u] = q.gt;
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

// exact_only != 0: every unit through the 64-bit exact kernel.  Otherwise fast kernel, then the
// exact kernel over the units the fast kernel could not certify (worklist of capacity wcap; the
// caller checks *wcount <= wcap afterwards and re-runs with exact_only on overflow).
// cut_chain >= 0: the sort-free staircase kernel (optics_rorr_cut.cu) instead of the fast sorting
// kernel; 0 = chains over the layer slots of a column, 1 = chains over cut_len consecutive columns.
void rt_launch_optics_rorr(const DevTables &tb, const DevChunk &ch, int *worklist, int *wcount,
                           int wcap, int exact_only, int cut_chain, int cut_len, cudaStream_t st) {
  const long long units = (long long)ch.ncol * tb.nw * ch.nslot_cap;
  // RORR unit ids run over columns rounded up to a multiple of 4 (column-grouped unit order)
  const long long runits = (long long)((ch.ncol + 3) / 4 * 4) * tb.nw * ch.nslot_cap;
  continua_kernel<<<(unsigned)std::min<long long>((units + 255) / 256, 148 * 32), 256, 0, st>>>(tb, ch);
  if (exact_only || !tb.rows_sorted) {
    rorr_launch_exact(tb, ch, nullptr, nullptr, 0, runits, st);
    return;
  }
  cudaMemsetAsync(wcount, 0, sizeof(int), st);
  if (cut_chain >= 0) rorr_launch_cut(tb, ch, cut_chain, cut_len, worklist, wcount, wcap, st);
  else rorr_launch_fast(tb, ch, worklist, wcount, wcap, runits, st);
  rorr_launch_exact(tb, ch, worklist, wcount, wcap, wcap, st);
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
