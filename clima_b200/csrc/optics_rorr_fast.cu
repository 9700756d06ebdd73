// optics_rorr_fast.cu -- instantiates the fast flavour of the optics + RORR kernel (rorr_impl.cuh).
// keys per lane at P = 16 and a private namespace: the two flavours use different lane mappings
#ifndef RORR_KPL16
#define RORR_KPL16 32
#endif
#include <cstdlib>
#define RORR_NS rorr_fast
#include "rorr_impl.cuh"

namespace {
template <int P, bool FULL>
void launch(const DevTables &tb, const DevChunk &ch, int *worklist, int *wcount, int wcap,
            long long units, cudaStream_t st) {
  using C = rorr_fast::Cfg<P>;
  auto kern = rorr_fast::optics_rorr_kernel<P, FULL, true>;
  // the opt-in above 48 KB of dynamic shared memory is per device: one flag per device ordinal
  static bool attr_set[64] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64 || !attr_set[dev]) {
    // on failure the launch below fails too and run_chunk reports it (cudaGetLastError)
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::BLOCK_BYTES + 65536) == cudaSuccess &&
        dev >= 0 && dev < 64)
      attr_set[dev] = true;
  }
  // One group of UPB units per block (NOT a persistent grid: the kernel is instruction-cache bound
  // and the warps of a block that start together share their instruction stream; measured r1n:
  // a persistent grid of 592 blocks doubles the no-instruction stalls)
  const long long nb = (units + C::UPB - 1) / C::UPB;
  if (nb <= 0) return;
  // experiment knob: extra dynamic shared memory lowers the blocks per SM (room for a concurrent kernel)
  static const size_t pad = [] { const char *e = getenv("CLIMA_RT_RORR_SMEM_PAD"); return e ? (size_t)atoi(e) : (size_t)0; }();
  kern<<<(unsigned)nb, 32 * RORR_WPB, C::BLOCK_BYTES + pad, st>>>(tb, ch, worklist, wcount, wcap);
}
}  // namespace

// units: number of work units to cover (all units of the chunk, or the worklist capacity).
void rorr_launch_fast(const DevTables &tb, const DevChunk &ch, int *worklist, int *wcount, int wcap,
                    long long units, cudaStream_t st) {
  const int G = tb.G;
  if (G <= 8) {
    if (G == 8) launch<8, true>(tb, ch, worklist, wcount, wcap, units, st);
    else launch<8, false>(tb, ch, worklist, wcount, wcap, units, st);
  } else if (G <= 16) {
    if (G == 16) launch<16, true>(tb, ch, worklist, wcount, wcap, units, st);
    else launch<16, false>(tb, ch, worklist, wcount, wcap, units, st);
  } else {
    if (G == 32) launch<32, true>(tb, ch, worklist, wcount, wcap, units, st);
    else launch<32, false>(tb, ch, worklist, wcount, wcap, units, st);
  }
}
