// rt_kernels.h -- host-callable launchers of the RT kernels (internal; the public surface is
// include/clima_rt.h).
#pragma once
#include <cuda_runtime.h>

#include "rt_device.cuh"

// Outputs of the two-stream + integration stages for one chunk.
struct DevFlux {
  // spectral, ground-first, [col][ll][level]
  double *fup_a_ir, *fdn_a_ir, *fup_a_sol, *fdn_a_sol, *amean_sol;
  // integrated [col][level]
  double *fup_n_ir, *fdn_n_ir, *fup_n_sol, *fdn_n_sol, *f_total;
  double *ISR, *OLR;       // [col]
  double *bpl;             // [col][ll_ir][nz+1] Planck, top-first
  double *scratch;         // two-stream forward-sweep storage, [layer][item][resident thread]
  size_t scratch_threads;  // resident threads the scratch was sized for
  int *status;
};

// prep.cu
void rt_launch_prep(const DevTables &tb, const DevChunk &ch, cudaStream_t st);
// counts the groups of 4 adjacent columns with unlike k-gas amounts into ch.dissim[0] (RORR unit order) and those
// with neighbours more than 25 % apart into ch.dissim[1] (staircase kernel or not)
void rt_launch_similarity(const DevTables &tb, const DevChunk &ch, cudaStream_t st);
// optics_rorr.cu
void rt_launch_optics_rorr(const DevTables &tb, const DevChunk &ch, int *worklist, int *wcount,
                           int wcap, int exact_only, int cut_chain, int cut_len, cudaStream_t st);
// twostream.cu.  `ocol_bcast`: every column reads the opacity of column 0 (frozen-opacity mode).
void rt_launch_planck(const DevTables &tb, const DevChunk &ch, const DevFlux &fx, cudaStream_t st);
void rt_launch_twostream_ir(const DevTables &tb, const DevParams &pr, const DevChunk &ch,
                            const DevFlux &fx, int ocol_bcast, int nblocks, cudaStream_t st);
void rt_launch_twostream_solar(const DevTables &tb, const DevParams &pr, const DevChunk &ch,
                               const DevFlux &fx, int ocol_bcast, int nblocks, cudaStream_t st);
void rt_launch_integrate(const DevTables &tb, const DevParams &pr, const DevChunk &ch,
                         const DevFlux &fx, int compute_solar, int sol_bcast, cudaStream_t st);
void rt_launch_tau_band(const DevTables &tb, const DevChunk &ch, int is_ir, double *out,
                        cudaStream_t st);
void rt_launch_expand_opacity(const DevTables &tb, const DevChunk &ch, int icol, double *tau,
                              double *w0, double *g, double *tband, cudaStream_t st);
void rt_twostream_occupancy(int G, int nzen, int *ir_blocks_per_sm, int *sol_blocks_per_sm);
int rt_twostream_block_threads();
int rt_twostream_items_per_layer();
// jacobian_delta.cu (N1: shared factorisation + delta solves)
size_t rt_jac_factor_doubles(const DevTables &tb, int nz);
void rt_launch_jac_factor(const DevTables &tb, const DevParams &pr, const DevChunk &ch, double *factor,
                          cudaStream_t st);
int rt_jac_splits(const DevTables &tb, int ncol, int num_sms);
void rt_launch_jac_delta(const DevTables &tb, const DevParams &pr, int nzm, int ncol, const double *Tm,
                         const double *Ts, const double *baseT, const double *factor, int *win, int *flag,
                         double *partial, int nsplit, cudaStream_t st);
void rt_launch_jac_reduce(const double *partial, const double *basef, double *out, int nzm, int ncol, int nsplit,
                          cudaStream_t st);
// fp64_peak.cu
double rt_measure_fp64_tflops(int device, cudaStream_t st);

// adiabat_profile.cu (N4: batched AdiabatClimate%make_profile)
#ifndef ADIABAT_MAXG
#define ADIABAT_MAXG 16  // gases
#endif
#ifndef ADIABAT_GROUP_MAX_COLS_PER_SM
#define ADIABAT_GROUP_MAX_COLS_PER_SM 64  // larger batches take the one-thread-per-column kernel
#endif
#define ADIABAT_MAXT 4   // temperature ranges of a heat-capacity polynomial
#ifndef ADIABAT_THREADS
#define ADIABAT_THREADS 32  // one warp per block: a batch spreads over the SMs
#endif
#ifndef ADIABAT_WARPS_PER_SM
#define ADIABAT_WARPS_PER_SM 4  // a warp takes more than one column only beyond this many warps per SM
#endif
struct AdiabatData {  // device copy of what make_profile reads from the AdiabatClimate object
  int ng, nz, double_radiative_grid;
  double mass[ADIABAT_MAXG];
  int model[ADIABAT_MAXG];   // 1 Shomate, 2 NASA-9
  int ntemps[ADIABAT_MAXG];
  double temps[ADIABAT_MAXG][ADIABAT_MAXT + 1];
  double coef[ADIABAT_MAXG][ADIABAT_MAXT][9];
  int has_sat[ADIABAT_MAXG];
  double sat[ADIABAT_MAXG][16];  // mu, T_ref, P_ref, T_triple, T_critical, a_v, b_v, a_s, b_s, a_c, b_c + 5 derived
  double RH[ADIABAT_MAXG];
  double planet_mass, planet_radius, P_top, T_trop, rtol, atol;
  double reference_pressure;  // <= 0: planet radius at the surface (clima_adiabat.f90:64)
  // particles (AdiabatClimate%set_particle_density_and_radii): np species, npl table levels = the given pressures
  // plus the two end points; device arrays part_x (npl) ascending log10 P, part_d / part_r (npl, np) log10 values
  int np, npl;
  const double *part_x, *part_d, *part_r;
};
struct AdiabatOut {  // device pointers
  double *P, *T, *z, *dz;            // (nz, ncol)
  double *f_i;                       // (nz, ng, ncol)
  double *f_i_surf, *N_surface;      // (ng, ncol)
  double *P_surf, *P_trop;           // (ncol)
  int *status, *nfcn;                // (ncol)
  double *T_r, *P_r, *dz_r, *dens_r, *Tsurf_r;  // radiative grid (nullptr: not wanted)
  double *pd_r, *rd_r;               // (nz_r, np, ncol) particle densities / radii (np > 0 only)
  const double *T_trop_in;           // (ncol) tropopause temperature per column, or nullptr: AdiabatData::T_trop
};
size_t rt_adiabat_work_doubles(int nz, int ncol);
void rt_launch_make_profile(const AdiabatData *a, int ng, int ncol, int num_sms, const double *T_surf,
                            const double *P_i_surf, const AdiabatOut &o, double *work, cudaStream_t st);

// the same launch with the per-column work arrays in a file-scope static __device__ array (no allocation anywhere);
// RT_NOALLOC_MAX_WORK_DOUBLES doubles = 16 MiB held by every GPU that loads the library
#ifndef RT_NOALLOC_MAX_WORK_DOUBLES
#define RT_NOALLOC_MAX_WORK_DOUBLES (2 * 1024 * 1024)
#endif
size_t rt_noalloc_max_work_doubles();
int rt_launch_make_profile_noalloc(const AdiabatData *a, int ng, int nz, int ncol, int num_sms, const double *T_surf,
                                   const double *P_i_surf, const AdiabatOut &o, cudaStream_t st, const char **why);
