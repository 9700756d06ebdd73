# cython/batch_methods.pyx -- the two methods the Cython layer gains (rows N3 of SURVEY.md section 8f),
# written against clima/cython/Radtran.pyx and clima/cython/AdiabatClimate.pyx:169-193.
# SOURCE ONLY: it compiles only inside the reference's scikit-build tree (needs the Fortran library).
#
# pxd additions (clima/cython/Radtran_pxd.pxd, AdiabatClimate_pxd.pxd):
#   cdef extern void radtran_radiate_batch_wrapper(Radtran *ptr, int *ncol, int *nz, int *ng,
#       double *T_surface, double *T, double *P, double *densities, double *dz, cbool *compute_solar,
#       double *fup_n_ir, double *fdn_n_ir, double *fup_n_sol, double *fdn_n_sol, double *f_total,
#       double *ISR, double *OLR, char *err)
#   cdef extern void adiabatclimate_toa_fluxes_batch_wrapper(AdiabatClimate *ptr, int *ncol, double *T_surf,
#       int *ng, double *P_i_surf, double *ISR, double *OLR, char *err)

# ---- in cdef class Radtran:
  def radiate_batch(self, ndarray[double, ndim=1] T_surface, ndarray[double, ndim=2] T,
                    ndarray[double, ndim=2] P, ndarray[double, ndim=3] densities,
                    ndarray[double, ndim=2] dz, cbool compute_solar = True):
    """`radiate` + `TOA_fluxes` for a batch of columns on the GPU (libclima_rt).

    Parameters
    ----------
    T_surface : ndarray[double,ndim=1]   (ncol) K
    T, P, dz : ndarray[double,ndim=2]    (nz,ncol) K, bar, cm -- Fortran ordered, ground first
    densities : ndarray[double,ndim=3]   (nz,ng,ncol) molecules/cm^3

    Returns
    -------
    dict with fup_n_ir, fdn_n_ir, fup_n_sol, fdn_n_sol, f_total (nz+1,ncol) [mW/m^2], ISR, OLR (ncol)
    """
    cdef int nz = T.shape[0], ncol = T.shape[1], ng = densities.shape[1]
    if not (T.flags.f_contiguous and P.flags.f_contiguous and densities.flags.f_contiguous and dz.flags.f_contiguous):
      raise ClimaException('"T", "P", "densities" and "dz" must be Fortran ordered')
    if T_surface.shape[0] != ncol or P.shape[0] != nz or P.shape[1] != ncol or dz.shape[0] != nz \
       or dz.shape[1] != ncol or densities.shape[0] != nz or densities.shape[2] != ncol:
      raise ClimaException('input arrays have inconsistent dimensions')
    cdef char err[ERR_LEN+1]
    names = ("fup_n_ir", "fdn_n_ir", "fup_n_sol", "fdn_n_sol", "f_total")
    out = {k: np.empty((nz+1, ncol), np.double, order="F") for k in names}
    cdef ndarray[double, ndim=2] a0 = out["fup_n_ir"], a1 = out["fdn_n_ir"], a2 = out["fup_n_sol"]
    cdef ndarray[double, ndim=2] a3 = out["fdn_n_sol"], a4 = out["f_total"]
    cdef ndarray[double, ndim=1] ISR = np.empty(ncol, np.double), OLR = np.empty(ncol, np.double)
    rad_pxd.radtran_radiate_batch_wrapper(self._ptr, &ncol, &nz, &ng, <double *>T_surface.data, <double *>T.data,
        <double *>P.data, <double *>densities.data, <double *>dz.data, &compute_solar,
        <double *>a0.data, <double *>a1.data, <double *>a2.data, <double *>a3.data, <double *>a4.data,
        <double *>ISR.data, <double *>OLR.data, err)
    if len(err.strip()) > 0:
      raise ClimaException(err.decode("utf-8").strip())
    out["ISR"] = ISR; out["OLR"] = OLR
    return out

# ---- in cdef class AdiabatClimate:
  def TOA_fluxes_batch(self, ndarray[double, ndim=1] T_surf, ndarray[double, ndim=2] P_i_surf):
    """`TOA_fluxes` for ncol (T_surf, P_i_surf) pairs: `make_profile` per column on the host, then one
    batched radiative-transfer call on the GPU (the loop of examples/EarlyMars.ipynb, cell 7).

    Parameters
    ----------
    T_surf : ndarray[double,ndim=1]     (ncol) surface temperatures (K)
    P_i_surf : ndarray[double,ndim=2]   (ng,ncol) surface pressures of each species (dynes/cm^2), Fortran ordered

    Returns
    -------
    tuple (ISR, OLR), each ndarray (ncol)
    """
    cdef int ng = P_i_surf.shape[0], ncol = P_i_surf.shape[1]
    if T_surf.shape[0] != ncol:
      raise ClimaException('"T_surf" and "P_i_surf" have inconsistent dimensions')
    if not P_i_surf.flags.f_contiguous:
      raise ClimaException('"P_i_surf" must be Fortran ordered')
    cdef char err[ERR_LEN+1]
    cdef ndarray[double, ndim=1] ISR = np.empty(ncol, np.double), OLR = np.empty(ncol, np.double)
    wa_pxd.adiabatclimate_toa_fluxes_batch_wrapper(self._ptr, &ncol, <double *>T_surf.data, &ng,
        <double *>P_i_surf.data, <double *>ISR.data, <double *>OLR.data, err)
    if len(err.strip()) > 0:
      raise ClimaException(err.decode("utf-8").strip())
    return ISR, OLR

  def TOA_fluxes_batch_device(self, ndarray[double, ndim=1] T_surf, ndarray[double, ndim=2] P_i_surf):
    """Same result, with `make_profile` itself on the GPU (row N4: fortran/clima_adiabat_cuda.f90,
    `AdiabatClimate%TOA_fluxes_batch_device`).  Not for objects with ocean solubility functions, particles,
    `reference_pressure > 0` or a temperature-dependent albedo: those raise and `TOA_fluxes_batch` applies.
    pxd addition: the wrapper below with the argument list of adiabatclimate_toa_fluxes_batch_wrapper."""
    cdef int ng = P_i_surf.shape[0], ncol = P_i_surf.shape[1]
    if T_surf.shape[0] != ncol:
      raise ClimaException('"T_surf" and "P_i_surf" have inconsistent dimensions')
    if not P_i_surf.flags.f_contiguous:
      raise ClimaException('"P_i_surf" must be Fortran ordered')
    cdef char err[ERR_LEN+1]
    cdef ndarray[double, ndim=1] ISR = np.empty(ncol, np.double), OLR = np.empty(ncol, np.double)
    wa_pxd.adiabatclimate_toa_fluxes_batch_device_wrapper(self._ptr, &ncol, <double *>T_surf.data, &ng,
        <double *>P_i_surf.data, <double *>ISR.data, <double *>OLR.data, err)
    if len(err.strip()) > 0:
      raise ClimaException(err.decode("utf-8").strip())
    return ISR, OLR
