/*
 * clima_rt.h -- C ABI of the B200-native radiative-transfer hot path of Clima.
 *
 * This is the drop-in boundary.  Every entry point below replaces (for a batch of
 * atmospheric columns) one piece of the reference's Fortran path and follows the
 * reference's own C-ABI conventions (clima/fortran/clima_c_api.f90:5,57-69 and
 * clima/fortran/Radtran.f90:77-107):
 *
 *   - every export is a plain C function returning void ("bind(c) subroutine");
 *   - the first argument of a method is the opaque handle, by value;
 *   - ALL scalars are passed by reference (int32_t*, double*, bool as int32_t*);
 *   - arrays are column-major (Fortran order) doubles, preceded by their dims;
 *   - the last argument is `char err[CLIMA_RT_ERR_LEN+1]`: "" on success, otherwise a
 *     NUL-terminated message (the Cython layer raises ClimaException when
 *     len(err.strip()) > 0, clima/cython/AdiabatClimate.pyx:190-192).
 *
 * There are no torch types, no C++ types and no CUDA types in any signature: host
 * pointers in, host pointers out (the `_device` variants take raw device pointers).
 * There is NO CPU fallback: every compute entry fails with an err message if no
 * CUDA device is usable.
 *
 * Layout conventions (identical to the reference):
 *   - inputs are ground-first (index 0 = bottom layer), pressures in bar, number
 *     densities in molecules/cm^3, dz in cm  (src/radtran/clima_radtran.f90:221-233);
 *   - batched arrays put the column index LAST (slowest): T(nz,ncol),
 *     densities(nz,ng,ncol), ... so one column's block is exactly the array the
 *     reference's single-column `radiate` takes;
 *   - outputs are ground-first, level index nz (0-based) = top of atmosphere
 *     (src/radtran/clima_radtran_radiate.f90:138-154).
 */
#ifndef CLIMA_RT_H
#define CLIMA_RT_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CLIMA_RT_ERR_LEN 1024 /* clima/fortran/clima_c_api.f90:5 (err_len) */

/* status bits written per column into clima_rt_out.status */
#define CLIMA_RT_STATUS_RADIUS_CLAMPED 1 /* particle radius outside table: the reference fails the
                                            whole call (clima_radtran_types.f90:972-976,773-776) */
#define CLIMA_RT_STATUS_NONFINITE 2      /* a non-finite flux was produced */

/* ---------------------------------------------------------------------------------
 * Read-only opacity tables.  Mirrors `type OpticalProperties` and its members
 * (src/radtran/clima_radtran_types.f90:23-141) after `create_OpticalProperties` has
 * rebinned everything onto the opacity grid
 * (src/radtran/clima_radtran_types_create.f90:272-645).  All indices are 1-based
 * species indices exactly as the Fortran holds them (sp_ind, p_ind, LH2O).
 * All pointers are HOST pointers; clima_rt_create copies them to the device.
 * ------------------------------------------------------------------------------- */
typedef struct clima_rt_tables {
  /* opacity grid: OpticalProperties%wavl (nm, ascending), freq = c/(wavl*1e-9) */
  int32_t nw;
  const double *wavl; /* (nw+1) */

  int32_t ng; /* Radtran%ng: gases in the model   (clima_radtran.f90:33) */
  int32_t np; /* Radtran%np: particles in the model (clima_radtran.f90:36) */

  /* k-distributions: type Ktable (clima_radtran_types.f90:23-38); all tables share
     `weights` within 1e-12 (clima_radtran_types_create.f90:379-389); Ksettings wbin/wbin_e/wxy
     are derived from them as in create_Ksettings (:191-224). */
  int32_t nk;
  int32_t ngauss;
  const double *weights;         /* (ngauss) of the first table */
  const int32_t *k_sp_ind;       /* (nk) */
  const int32_t *k_npress;       /* (nk) */
  const int32_t *k_ntemp;        /* (nk) */
  const double *const *k_log10P; /* nk pointers, each (npress) log10(bar), strictly increasing */
  const double *const *k_temp;   /* nk pointers, each (ntemp) K, strictly increasing */
  const double *const *k_log10k; /* nk pointers, each (ngauss,npress,ntemp,nw) column-major,
                                    log10(cm^2/molecule)  (types_create.f90:1291-1358) */

  /* CIA: type Xsection with xs_type=CIAXsection (clima_radtran_types.f90:44-55).
     dim 1: data = log10 xs (ntemp,nw) column-major, interpolated linearly in clamped T;
     dim 0: data = xs_0d (nw). */
  int32_t ncia;
  const int32_t *cia_sp_ind; /* (2,ncia) column-major */
  const int32_t *cia_dim;    /* (ncia) 0 or 1 */
  const int32_t *cia_ntemp;  /* (ncia) (ignored for dim 0) */
  const double *const *cia_temp;
  const double *const *cia_data;

  /* Rayleigh: xs_0d per species (types_create.f90:1048-1088) */
  int32_t nray;
  const int32_t *ray_sp_ind; /* (nray) */
  const double *ray_xs;      /* (nw,nray) column-major */

  /* photolysis / absorption cross sections: same layout as CIA with one species each */
  int32_t npxs;
  const int32_t *pxs_sp_ind; /* (npxs) */
  const int32_t *pxs_dim;
  const int32_t *pxs_ntemp;
  const double *const *pxs_temp;
  const double *const *pxs_data;

  /* water continuum: type WaterContinuum (clima_radtran_types.f90:68-77) */
  int32_t has_cont;
  int32_t cont_LH2O;
  int32_t cont_ntemp;
  const double *cont_temp;          /* (ntemp) */
  const double *cont_log10_H2O;     /* (ntemp,nw) column-major */
  const double *cont_log10_foreign; /* (ntemp,nw) */

  /* particles: type ParticleXsection (clima_radtran_types.f90:57-66) */
  int32_t npart;
  const int32_t *part_p_ind; /* (npart) */
  const int32_t *part_nrad;  /* (npart) */
  const double *const *part_radii; /* each (nrad) cm, strictly increasing */
  const double *const *part_w0;    /* each (nrad,nw) */
  const double *const *part_qext;  /* each (nrad,nw) */
  const double *const *part_gt;    /* each (nrad,nw) */

  /* channels: type RTChannel (clima_radtran_types.f90:263-269): 1-based inclusive bin ranges
     into the opacity grid (types_create.f90:250-268). */
  int32_t ir_ind_start, ir_ind_end;
  int32_t sol_ind_start, sol_ind_end;

  /* Radtran%photons_sol (nw_sol) mW/m^2/Hz (clima_radtran.f90:63-65) */
  const double *photons_sol;
} clima_rt_tables;

/* ---------------------------------------------------------------------------------
 * Run-time knobs: the public members of `type Radtran` (clima_radtran.f90:51-68).
 * ------------------------------------------------------------------------------- */
typedef struct clima_rt_params {
  double diurnal_fac;         /* default 0.5 */
  double photon_scale_factor; /* default 1.0 */
  double ir_tau_min;          /* default 1e-6 */
  int32_t has_hard_surface;   /* default 1 */
  int32_t nzenith;
  const double *zenith_u;       /* (nzenith) cos(zenith) */
  const double *zenith_weights; /* (nzenith) */
  const double *surface_albedo;     /* (nw_sol) */
  const double *surface_emissivity; /* (nw_ir) */
} clima_rt_params;

/* ---------------------------------------------------------------------------------
 * Result pointers: `type ClimaRadtranWrk` for both channels + f_total
 * (clima_radtran.f90:11-25,70-72) + TOA_fluxes' ISR/OLR (:339-340), per column.
 * Mandatory: fup_n_*, fdn_n_*, f_total, ISR, OLR.  Any spectral pointer may be NULL.
 * ------------------------------------------------------------------------------- */
typedef struct clima_rt_out {
  double *fup_n_ir;  /* (nz+1,ncol) mW/m^2 */
  double *fdn_n_ir;  /* (nz+1,ncol) */
  double *fup_n_sol; /* (nz+1,ncol) */
  double *fdn_n_sol; /* (nz+1,ncol) */
  double *f_total;   /* (nz+1,ncol) */
  double *ISR;       /* (ncol) */
  double *OLR;       /* (ncol) */
  /* optional spectral outputs (NULL = not wanted) */
  double *fup_a_ir;     /* (nz+1,nw_ir,ncol) mW/m^2/Hz */
  double *fdn_a_ir;     /* (nz+1,nw_ir,ncol) */
  double *tau_band_ir;  /* (nz,nw_ir,ncol) */
  double *fup_a_sol;    /* (nz+1,nw_sol,ncol) */
  double *fdn_a_sol;    /* (nz+1,nw_sol,ncol) */
  double *amean_sol;    /* (nz+1,nw_sol,ncol) photons/cm^2/s */
  double *tau_band_sol; /* (nz,nw_sol,ncol) */
  int32_t *status;      /* (ncol) CLIMA_RT_STATUS_* bits, may be NULL */
} clima_rt_out;

/* Create / destroy.  Replaces the table-holding part of `create_Radtran_2`
   (clima_radtran.f90:128-219): uploads the tables to device `*device` and allocates nothing
   per-column yet.  `*handle` receives an opaque pointer (like allocate_adiabatclimate,
   clima/fortran/AdiabatClimate.f90:7-22). */
void clima_rt_create(const clima_rt_tables *tables, const clima_rt_params *params,
                     const int32_t *device, void **handle, char *err);
void clima_rt_destroy(void *handle);

/* Replace the run-time knobs (setters of clima/fortran/Radtran.f90:130-300). */
void clima_rt_set_params(void *handle, const clima_rt_params *params, char *err);

/* Custom optical properties already interpolated onto the opacity grid, i.e. the state that
   OpticalProperties_set_custom_optical_properties leaves in dtau_dz_interp/w0_interp/g0_interp
   (clima_radtran_types.f90:504-531): log10P (nP, ascending, log10 dyn/cm^2) and (nP,nw) tables.
   clima_rt_unset_custom_optical_properties restores the `tiny(0d0)` behaviour (:558-563). */
void clima_rt_set_custom_optical_properties(void *handle, const int32_t *nP, const double *log10P,
                                            const double *dtau_dz, const double *w0,
                                            const double *g0, char *err);
void clima_rt_unset_custom_optical_properties(void *handle);

/* The batched-column entry point: `Radtran%radiate` + `Radtran%TOA_fluxes`
   (clima_radtran.f90:221-342) for ncol independent columns.
     T_surface(ncol), T(nz,ncol), P(nz,ncol) [bar], densities(nz,ng,ncol), dz(nz,ncol),
     pdensities(nz,np,ncol) / radii(nz,np,ncol) (both NULL or both given; mandatory if np>0,
     clima_radtran.f90:426-436).
   compute_opacity=0 reuses the opacity retained in the handle: either one set per column of the
   previous call (same ncol, single chunk) or ONE column's set broadcast to all columns (the
   frozen-opacity Jacobian mode of src/adiabat/clima_adiabat_solve.f90:798-820).
   compute_solar=0 reuses the stale solar fluxes in f_total (clima_radtran.f90:286-289).
   All data pointers are HOST pointers. */
void clima_rt_radiate_batch(void *handle, const int32_t *ncol, const int32_t *nz, const int32_t *ng,
                            const int32_t *np, const double *T_surface, const double *T,
                            const double *P, const double *densities, const double *dz,
                            const double *pdensities, const double *radii,
                            const int32_t *compute_solar, const int32_t *compute_opacity,
                            const clima_rt_out *out, char *err);

/* Same, but every data pointer (inputs and `out` members) is a DEVICE pointer on the handle's
   device; work is enqueued on the handle's stream and the call returns after a stream sync. */
void clima_rt_radiate_batch_device(void *handle, const int32_t *ncol, const int32_t *nz,
                                   const int32_t *ng, const int32_t *np, const double *T_surface,
                                   const double *T, const double *P, const double *densities,
                                   const double *dz, const double *pdensities, const double *radii,
                                   const int32_t *compute_solar, const int32_t *compute_opacity,
                                   const clima_rt_out *out, char *err);

/* ---- multi-GPU.  Columns are independent: every rank (one process or thread per GPU,
   each with its own handle and a replica of the tables) computes a contiguous block of columns and
   ONE all-gather returns every rank's flux profiles to every rank.  The communicator is NCCL's
   (bound at run time with dlopen: no link-time dependency, a single-GPU caller needs no NCCL).
     clima_rt_comm_unique_id   rank 0 obtains the id (ncclGetUniqueId) and hands the 128 bytes to the other
                               ranks by its own means (MPI_Bcast, a file, torch.distributed ...)
     clima_rt_comm_init        collective: ncclCommInitRank on the handle's device
     clima_rt_comm_destroy     also called by clima_rt_destroy
   The reference has no multi-device path; the caller these replace is the loop over columns of a
   sweep (examples/EarlyMars.ipynb cell 7) split over processes. */
#define CLIMA_RT_COMM_ID_BYTES 128
void clima_rt_comm_unique_id(char *id /* [CLIMA_RT_COMM_ID_BYTES] */, char *err);
void clima_rt_comm_init(void *handle, const int32_t *nranks, const int32_t *rank,
                        const char *id /* [CLIMA_RT_COMM_ID_BYTES] */, char *err);
void clima_rt_comm_destroy(void *handle);

/* ncclAllGather of `*count` doubles per rank between DEVICE buffers (recv holds nranks * count), on
   the handle's stream: ordered after the kernels of a preceding clima_rt_radiate_batch_device, no
   host synchronisation.  The building block of clima_rt_radiate_batch_sharded; bench.py times it. */
void clima_rt_allgather_device(void *handle, const double *send, double *recv, const int64_t *count,
                               char *err);

/* clima_rt_radiate_batch for this rank's `*ncol` columns (HOST inputs, 0 <= ncol <= ncol_max; ncol_max
   is the same on every rank) followed by the all-gather: every rank receives
     fup_n_ir, fdn_n_ir, fup_n_sol, fdn_n_sol, f_total (nz+1, ncol_max, nranks), ISR, OLR (ncol_max, nranks)
   in HOST memory; rank r's columns are [.., 0 .. ncol_r) of slab r, the padding columns are zero.
   Spectral pointers of out_all are ignored (727 KB per column stay on the rank that computed them). */
void clima_rt_radiate_batch_sharded(void *handle, const int32_t *ncol, const int32_t *ncol_max,
                                    const int32_t *nz, const int32_t *ng, const int32_t *np,
                                    const double *T_surface, const double *T, const double *P,
                                    const double *densities, const double *dz, const double *pdensities,
                                    const double *radii, const int32_t *compute_solar,
                                    const int32_t *compute_opacity, const clima_rt_out *out_all, char *err);

/* N1: the temperature-perturbation loop of AdiabatClimate_jacobian_from_base
   (src/adiabat/clima_adiabat_solve.f90:798-820) as ONE call.  The handle must hold the opacity and
   the solar fluxes of the BASE column on the doubled radiative grid (a previous
   clima_rt_radiate_batch with ncol=1, nz_r = 2*nz+2, compute_opacity=1, compute_solar=1, i.e. the
   objective evaluation that precedes the Jacobian in the reference, :757-761).  Inputs are on the
   MODEL grid: T_surface(ncol), T(nz,ncol), one perturbed temperature profile per column.  On the
   device: AdiabatClimate_copy_atm_to_radiative_grid (src/adiabat/clima_adiabat.f90:729-773,
   temperatures only: each layer twice + two ghost copies of the top layer), then `radiate` with
   compute_opacity=.false., compute_solar=.false. for every column (opacity and stale solar fluxes
   broadcast), then f_total(i) = rad%f_total(2*i-1) (:731-733).
   Output: f_total(nz+1,ncol) [mW/m^2], WITHOUT surface_heat_flow (the caller adds it, :734).
   All pointers are HOST pointers. */
void clima_rt_jacobian_batch(void *handle, const int32_t *ncol, const int32_t *nz,
                             const double *T_surface, const double *T, double *f_total, char *err);

/* ---- N4: `AdiabatClimate%make_profile` for a batch of (T_surf, P_i_surf) pairs, built on
   the device, and the batched `AdiabatClimate%TOA_fluxes` on top of it.

   clima_rt_adiabat is what make_profile reads from the AdiabatClimate object
   (src/adiabat/clima_adiabat.f90:13-135): the gases of species.yaml in the order of the radiative
   tables (mass, heat-capacity polynomial, optional LinearLatentHeat saturation data:
   src/clima_types.f90 `Gas`, src/clima_saturationdata.f90:16-28), RH, P_top, T_trop, the planet,
   rtol / atol and the number of model layers.  Not carried: ocean solubility functions (host
   callbacks that only fill N_ocean: the caller runs them on N_surface and f_i_surf * P_surf, as
   clima_b200/adiabat.py does), CustomDrySpecies.  Particle profiles:
   clima_rt_adiabat_set_particles below. */
#define CLIMA_RT_ADIABAT_MAX_GASES 16
#define CLIMA_RT_ADIABAT_MAX_TEMPS 4
typedef struct clima_rt_adiabat {
  int32_t ng;                    /* gases, = the ng of the radiative tables */
  int32_t nz;                    /* model layers (AdiabatClimate%nz) */
  int32_t double_radiative_grid; /* default 1: nz_r = 2 nz + 2 (clima_adiabat.f90:729-762) */
  const double *mass;            /* (ng) g/mol */
  const int32_t *thermo_model;   /* (ng) 1 = Shomate, 2 = NASA-9 (clima_eqns.f90:82-103) */
  const int32_t *thermo_ntemps;  /* (ng) temperature ranges, <= CLIMA_RT_ADIABAT_MAX_TEMPS */
  const double *thermo_temps;    /* (MAX_TEMPS + 1, ng) range bounds [K] */
  const double *thermo_data;     /* (9, MAX_TEMPS, ng) polynomial coefficients (7 used by Shomate) */
  const int32_t *has_sat;        /* (ng) 1 when the gas has saturation data */
  const double *sat;             /* (11, ng) mu, T-ref, P-ref, T-triple, T-critical, a_v, b_v, a_s, b_s, a_c, b_c */
  const double *RH;              /* (ng) relative humidity, default 1 */
  double planet_mass, planet_radius; /* g, cm */
  double P_top;                  /* dynes/cm^2, default 1 */
  double T_trop;                 /* K, default 180 */
  double rtol, atol;             /* defaults 1e-9, 1e-12 (clima_adiabat.f90:133-135) */
  double reference_pressure;     /* dynes/cm^2; <= 0 (default -1): the planet radius is at the surface, else at this
                                    pressure (clima_adiabat.f90:58-64, clima_adiabat_altitude.f90:97-169) */
} clima_rt_adiabat;

/* What make_profile leaves in the object, per column (HOST pointers; any may be NULL): */
typedef struct clima_rt_profile_out {
  double *P;         /* (nz,ncol) dynes/cm^2, layer centres */
  double *T;         /* (nz,ncol) K */
  double *z;         /* (nz,ncol) cm */
  double *dz;        /* (nz,ncol) cm */
  double *f_i;       /* (nz,ng,ncol) mixing ratios */
  double *f_i_surf;  /* (ng,ncol) */
  double *N_surface; /* (ng,ncol) mol/cm^2 condensed at the surface */
  double *P_surf;    /* (ncol) dynes/cm^2 */
  double *P_trop;    /* (ncol) dynes/cm^2, negative when no tropopause was found */
  int32_t *status;   /* (ncol) CLIMA_RT_PROFILE_* */
  int32_t *nfcn;     /* (ncol) right-hand-side evaluations of the adiabat integration */
} clima_rt_profile_out;
/* per-column status: the `err` strings of make_profile / compute_altitude */
#define CLIMA_RT_PROFILE_OK 0
#define CLIMA_RT_PROFILE_TSURF_BELOW_TTROP 1   /* clima_adiabat_general.f90:124-127 */
#define CLIMA_RT_PROFILE_NEGATIVE_PRESSURE 2   /* :128-131 */
#define CLIMA_RT_PROFILE_PTOP_ABOVE_SURFACE 3  /* :249-252 */
#define CLIMA_RT_PROFILE_DOP853_FAILED 4       /* :306-310 */
#define CLIMA_RT_PROFILE_BRENT_FAILED 5        /* :468-471 */
#define CLIMA_RT_PROFILE_HEAT_CAPACITY 6       /* :618-621 */
#define CLIMA_RT_PROFILE_NEGATIVE_ALTITUDE 7   /* :266-270 */
#define CLIMA_RT_PROFILE_ALTITUDE_FAILED 8     /* clima_adiabat_altitude.f90:215-220 */
#define CLIMA_RT_PROFILE_PRESSURE_GRID 9       /* clima_adiabat_altitude.f90:52-55 */
#define CLIMA_RT_PROFILE_NEGATIVE_TTROP 10     /* clima_adiabat_general.f90:136-139 */
#define CLIMA_RT_PROFILE_REFERENCE_PRESSURE 11 /* clima_adiabat_altitude.f90:97-105 */

/* Give the handle the adiabat description (copied to the device; call again after changing RH, T_trop, ...:
   the setters of clima/fortran/AdiabatClimate.f90). */
void clima_rt_adiabat_set(void *handle, const clima_rt_adiabat *adiabat, char *err);

/* `AdiabatClimate%set_particle_density_and_radii` (src/adiabat/clima_adiabat.f90:1047-1123) for handles whose tables
   have particles: P(n) [dynes/cm^2, strictly decreasing], pdensities(n,np) [particles/cm^3], pradii(n,np) [cm], HOST
   pointers, same checks and messages.  make_profile then interpolates both onto its pressure grid like
   `interpolate_particles` (:1022-1044: linear in log10 P of the log10 values, end values held) and the batched
   TOA_fluxes passes them to the RT.  Without this call the constructor's default applies (:387-389: no particles,
   radius 1 micron). */
void clima_rt_adiabat_set_particles(void *handle, const int32_t *n, const double *P, const double *pdensities,
                                    const double *pradii, char *err);

/* One tropopause temperature PER COLUMN for the following batch calls (which must then have ncol = *n), instead of
   clima_rt_adiabat::T_trop; *n = 0 returns to the single value.  The reference sets self%T_trop before each
   make_profile of its solvers when solve_for_T_trop is on (src/adiabat/clima_adiabat.f90:926-1000): in a batch every
   column has its own iterate. */
void clima_rt_adiabat_set_column_T_trop(void *handle, const int32_t *n, const double *T_trop, char *err);

/* `AdiabatClimate%make_profile` (src/adiabat/clima_adiabat.f90:401-472) for ncol columns:
   T_surf(ncol) [K], P_i_surf(ng,ncol) [dynes/cm^2], HOST pointers.  A column that fails gets its status
   code; `err` names the first such column with the reference's message, the others are still computed. */
void clima_rt_make_profile_batch(void *handle, const int32_t *ncol, const double *T_surf,
                                 const double *P_i_surf, const clima_rt_profile_out *profile, char *err);

/* `AdiabatClimate%TOA_fluxes` (src/adiabat/clima_adiabat.f90:776-799) for ncol columns in ONE call:
   make_profile on the device, copy_atm_to_radiative_grid written straight into the arrays the RT kernels
   read, then the radiate path.  `profile` may be NULL.  `out` as for clima_rt_radiate_batch with
   nz_r = 2 nz + 2 (nz when double_radiative_grid = 0) levels.  Any failed column fails the call (err set,
   `out` untouched), as the reference's loop would stop there. */
void clima_rt_adiabat_toa_fluxes_batch(void *handle, const int32_t *ncol, const double *T_surf,
                                       const double *P_i_surf, const clima_rt_profile_out *profile,
                                       const clima_rt_out *out, char *err);

/* clima_rt_make_profile_batch for callers that may not allocate (code under CUDA graph capture, tight serving
   loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
   Same kernel and same results as clima_rt_make_profile_batch, but
     * every pointer is a DEVICE pointer of the caller on the handle's device: T_surf(ncol), P_i_surf(ng,ncol)
       and ALL eleven members of `profile` (none may be NULL; status / nfcn are int32 (ncol));
     * the kernel's per-column work arrays live in a file-scope static __device__ array of
       CLIMA_RT_NOALLOC_MAX_WORK_DOUBLES doubles (16 MiB, held by every GPU that loads the library until the
       process exits) instead of a cudaMalloc'ed buffer; a call needs ncol * (7 nz + 5) of them (nz = 50: at
       most 5907 columns) and is rejected with a clear error beyond that;
     * the call itself issues no cudaMalloc / cudaFree / event creation: the handle and clima_rt_adiabat_set
       (which do allocate) are set-up and come first;
     * per-column failures are left in `profile->status` (CLIMA_RT_PROFILE_*) for the caller to read; `err` is
       set only for argument, size and launch errors;
     * a per-column tropopause temperature (clima_rt_adiabat_set_column_T_trop) applies as in the other calls.
   The work is enqueued on the handle's stream and the call returns after a stream synchronisation.  Every call
   (from every handle on that GPU) shares the one array, which keeps the last call's contents -- nothing in it
   has to start at zero, each column writes its work arrays before reading them -- so calls that OVERLAP on the
   GPU (two handles, two host threads) race; serialise them. */
#define CLIMA_RT_NOALLOC_MAX_WORK_DOUBLES (2 * 1024 * 1024)
void clima_rt_make_profile_batch_noalloc(void *handle, const int32_t *ncol, const double *T_surf,
                                         const double *P_i_surf, const clima_rt_profile_out *profile, char *err);

/* N4 bookkeeping: [0] device time (ms, CUDA events on the handle's stream) of the last make_profile kernel,
   [1] make_profile launches since the handle was created. */
void clima_rt_get_adiabat_stats(void *handle, double *stats2);

/* Copy the retained opacity of column `*icol` (0-based) of the last chunk to the host:
   tau(nz,ngauss,nw), w0(nz,ngauss,nw), g(nz,nw), tau_band(nz,nw), all TOP-first exactly like
   `type OpticalPropertiesResult` (clima_radtran_types.f90:242-247).  Test/debug hook. */
void clima_rt_get_opacity(void *handle, const int32_t *icol, double *tau, double *w0, double *g,
                          double *tau_band, char *err);

/* Introspection used by the bench: number of kernel launches issued by the handle so far, the
   per-stage device time (ms, CUDA events on the handle's stream) of the last batch call
   [0]=optics+RORR kernel [1]=Planck + two-stream IR [2]=two-stream solar [3]=integration
   [4]=whole call (kernels only, without the host<->device copies) [5]=prep kernel. */
void clima_rt_get_launch_count(void *handle, int64_t *count);
void clima_rt_get_stage_ms(void *handle, double *ms6);

/* RORR bookkeeping since the handle was created: [0] work units (column x bin x unique layer)
   processed, [1] units that went through the exact 64-bit sort kernel (all of them when the
   k-tables are not ascending in g or CLIMA_RT_RORR=exact is set; otherwise only those the 32-bit
   fast kernel could not certify), [2] worklist overflows (whole chunk redone exactly). */
void clima_rt_get_rorr_stats(void *handle, int64_t *stats3);

/* Bookkeeping of the sort-free (staircase) RORR kernel since the chunk buffers were allocated:
   [0] warp-merges (one merge of the 32/P units of a warp), [1] warp-passes over the rows (1 per merge
   when the predecessor's staircases are still the right ones), [2] restarts from bisection,
   [3] units handed to the exact sorting kernel, [4] cold starts (warp-merges at the head of a chain),
   [5] RORR kernel of the last chunk: 0 staircase/slot chains, 1 staircase/column chains, 2 sorting
   kernels, 3 exact sorting kernel only.  [6] warp-merges that took the closed form (rows or columns
   of a_i + b_j do not interleave).  [7] clima_rt_jacobian_batch calls served by the delta path
   (shared factorisation, jacobian_delta.cu). */
void clima_rt_get_rorr_cut_stats(void *handle, int64_t *stats8);

/* The CUDA stream (a cudaStream_t, returned as an opaque pointer) every kernel of the handle is
   launched on, so that a caller can bracket calls with its own CUDA events on the launching
   stream or order its own work (e.g. an NCCL all-gather of the flux profiles) after them. */
void clima_rt_get_stream(void *handle, void **stream);

/* Measured FP64 issue roof of device `*device`: runs a dependent-chain-free DFMA kernel and
   returns TFLOP/s (FMA = 2 flops).  Used as the roofline denominator. */
void clima_rt_measure_fp64_peak(const int32_t *device, double *tflops, char *err);

/* Library version string, NUL terminated (cf. clima_version_get, clima_c_api.f90:15-19). */
void clima_rt_version_get(char *version_c /* [101] */);

#ifdef __cplusplus
}
#endif
#endif /* CLIMA_RT_H */
