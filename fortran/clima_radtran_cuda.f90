!> clima_radtran_cuda -- iso_c_binding shim between Clima's Fortran `Radtran` type and the CUDA
!> C-ABI library `libclima_rt.so` (include/clima_rt.h).
!>
!> SOURCE ONLY in this repository: neither the build container nor the GPU box has a Fortran
!> compiler (SURVEY.md F2), so this file is written against the Fortran 2008 standard and the
!> reference's own conventions (clima/fortran/clima_c_api.f90:5,57-69) but has not been compiled
!> here.  INTEGRATION.md shows where it plugs into src/radtran/clima_radtran.f90.
!>
!> What it adds to the reference: a batched-column entry point
!>     call radtran_radiate_batch(rc, T_surface, T, P, densities, dz, ..., res, err)
!> next to the unchanged `rad%radiate` / `rad%TOA_fluxes`.  There is no CPU fallback: if the CUDA
!> library cannot create a handle, `err` is allocated.
module clima_radtran_cuda
  use iso_c_binding
  implicit none
  private

  public :: RadtranCuda, RadtranBatchResult
  public :: radtran_cuda_create, radtran_cuda_destroy, radtran_cuda_sync_params
  public :: radtran_radiate_batch, radtran_jacobian_batch, pack_params
  public :: c_tables, c_params, c_out

  integer, parameter :: dp = c_double
  integer, parameter :: err_len = 1024 !! clima/fortran/clima_c_api.f90:5

  !> struct clima_rt_tables (include/clima_rt.h): member order and types must match exactly.
  type, bind(c) :: c_tables
    integer(c_int32_t) :: nw
    type(c_ptr) :: wavl
    integer(c_int32_t) :: ng, np
    integer(c_int32_t) :: nk, ngauss
    type(c_ptr) :: weights, k_sp_ind, k_npress, k_ntemp, k_log10P, k_temp, k_log10k
    integer(c_int32_t) :: ncia
    type(c_ptr) :: cia_sp_ind, cia_dim, cia_ntemp, cia_temp, cia_data
    integer(c_int32_t) :: nray
    type(c_ptr) :: ray_sp_ind, ray_xs
    integer(c_int32_t) :: npxs
    type(c_ptr) :: pxs_sp_ind, pxs_dim, pxs_ntemp, pxs_temp, pxs_data
    integer(c_int32_t) :: has_cont, cont_LH2O, cont_ntemp
    type(c_ptr) :: cont_temp, cont_log10_H2O, cont_log10_foreign
    integer(c_int32_t) :: npart
    type(c_ptr) :: part_p_ind, part_nrad, part_radii, part_w0, part_qext, part_gt
    integer(c_int32_t) :: ir_ind_start, ir_ind_end, sol_ind_start, sol_ind_end
    type(c_ptr) :: photons_sol
  end type

  !> struct clima_rt_params: the public knobs of `type Radtran` (clima_radtran.f90:51-68)
  type, bind(c) :: c_params
    real(c_double) :: diurnal_fac, photon_scale_factor, ir_tau_min
    integer(c_int32_t) :: has_hard_surface, nzenith
    type(c_ptr) :: zenith_u, zenith_weights, surface_albedo, surface_emissivity
  end type

  !> struct clima_rt_out
  type, bind(c) :: c_out
    type(c_ptr) :: fup_n_ir, fdn_n_ir, fup_n_sol, fdn_n_sol, f_total, ISR, OLR
    type(c_ptr) :: fup_a_ir, fdn_a_ir, tau_band_ir
    type(c_ptr) :: fup_a_sol, fdn_a_sol, amean_sol, tau_band_sol
    type(c_ptr) :: status
  end type

  !> Batched counterpart of `ClimaRadtranWrk` x 2 + f_total (clima_radtran.f90:11-25,70-72):
  !> every array gains a trailing column dimension.
  type :: RadtranBatchResult
    real(dp), allocatable :: fup_n_ir(:,:), fdn_n_ir(:,:)   !! (nz+1,ncol) mW/m^2
    real(dp), allocatable :: fup_n_sol(:,:), fdn_n_sol(:,:) !! (nz+1,ncol)
    real(dp), allocatable :: f_total(:,:)                   !! (nz+1,ncol)
    real(dp), allocatable :: ISR(:), OLR(:)                 !! (ncol)
    integer(c_int32_t), allocatable :: status(:)            !! (ncol)
  end type

  !> Owns the opaque CUDA handle (like the c_ptr of allocate_adiabatclimate,
  !> clima/fortran/AdiabatClimate.f90:7-22).
  type :: RadtranCuda
    type(c_ptr) :: handle = c_null_ptr
    integer :: ng = 0, np = 0
  end type

  interface
    subroutine clima_rt_create(tables, params, device, handle, err) bind(c, name='clima_rt_create')
      import :: c_tables, c_params, c_int32_t, c_ptr, c_char
      type(c_tables), intent(in) :: tables
      type(c_params), intent(in) :: params
      integer(c_int32_t), intent(in) :: device
      type(c_ptr), intent(out) :: handle
      character(kind=c_char), intent(out) :: err(*)
    end subroutine
    subroutine clima_rt_destroy(handle) bind(c, name='clima_rt_destroy')
      import :: c_ptr
      type(c_ptr), value :: handle
    end subroutine
    subroutine clima_rt_set_params(handle, params, err) bind(c, name='clima_rt_set_params')
      import :: c_params, c_ptr, c_char
      type(c_ptr), value :: handle
      type(c_params), intent(in) :: params
      character(kind=c_char), intent(out) :: err(*)
    end subroutine
    subroutine clima_rt_radiate_batch(handle, ncol, nz, ng, np, T_surface, T, P, densities, dz, &
                                      pdensities, radii, compute_solar, compute_opacity, res, err) &
                                      bind(c, name='clima_rt_radiate_batch')
      import :: c_out, c_ptr, c_int32_t, c_double, c_char
      type(c_ptr), value :: handle
      integer(c_int32_t), intent(in) :: ncol, nz, ng, np
      real(c_double), intent(in) :: T_surface(*), T(*), P(*), densities(*), dz(*)
      type(c_ptr), value :: pdensities, radii !! NULL when the model has no particles
      integer(c_int32_t), intent(in) :: compute_solar, compute_opacity
      type(c_out), intent(in) :: res
      character(kind=c_char), intent(out) :: err(*)
    end subroutine
    subroutine clima_rt_jacobian_batch(handle, ncol, nz, T_surface, T, f_total, err) &
                                       bind(c, name='clima_rt_jacobian_batch')
      import :: c_ptr, c_int32_t, c_double, c_char
      type(c_ptr), value :: handle
      integer(c_int32_t), intent(in) :: ncol, nz
      real(c_double), intent(in) :: T_surface(*), T(*)
      real(c_double), intent(out) :: f_total(*)
      character(kind=c_char), intent(out) :: err(*)
    end subroutine
  end interface

contains

  subroutine copy_err(err_c, err)
    character(kind=c_char), intent(in) :: err_c(err_len+1)
    character(:), allocatable, intent(out) :: err
    integer :: i, n
    n = 0
    do i = 1,err_len
      if (err_c(i) == c_null_char) exit
      n = n + 1
    enddo
    if (n == 0) return  ! "" = success; err stays unallocated like every Clima routine
    allocate(character(n) :: err)
    do i = 1,n
      err(i:i) = err_c(i)
    enddo
  end subroutine

  !> Fills `c_params` from the public members of a Radtran object.  The caller passes the member
  !> arrays themselves (they must be `target`); nothing is copied on the Fortran side.
  subroutine pack_params(diurnal_fac, photon_scale_factor, ir_tau_min, has_hard_surface, &
                         zenith_u, zenith_weights, surface_albedo, surface_emissivity, pr)
    real(dp), intent(in) :: diurnal_fac, photon_scale_factor, ir_tau_min
    logical, intent(in) :: has_hard_surface
    real(dp), target, intent(in) :: zenith_u(:), zenith_weights(:)
    real(dp), target, intent(in) :: surface_albedo(:), surface_emissivity(:)
    type(c_params), intent(out) :: pr
    pr%diurnal_fac = diurnal_fac
    pr%photon_scale_factor = photon_scale_factor
    pr%ir_tau_min = ir_tau_min
    pr%has_hard_surface = merge(1_c_int32_t, 0_c_int32_t, has_hard_surface)
    pr%nzenith = int(size(zenith_u), c_int32_t)
    pr%zenith_u = c_loc(zenith_u)
    pr%zenith_weights = c_loc(zenith_weights)
    pr%surface_albedo = c_loc(surface_albedo)
    pr%surface_emissivity = c_loc(surface_emissivity)
  end subroutine

  !> Creates the CUDA handle from an already packed table descriptor (see
  !> `Radtran_pack_tables` in INTEGRATION.md, which lives next to the private table types).
  subroutine radtran_cuda_create(rc, tables, params, device, err)
    type(RadtranCuda), intent(inout) :: rc
    type(c_tables), intent(in) :: tables
    type(c_params), intent(in) :: params
    integer, intent(in) :: device
    character(:), allocatable, intent(out) :: err
    character(kind=c_char) :: err_c(err_len+1)
    call clima_rt_create(tables, params, int(device, c_int32_t), rc%handle, err_c)
    call copy_err(err_c, err)
    rc%ng = tables%ng
    rc%np = tables%np
  end subroutine

  subroutine radtran_cuda_destroy(rc)
    type(RadtranCuda), intent(inout) :: rc
    if (c_associated(rc%handle)) call clima_rt_destroy(rc%handle)
    rc%handle = c_null_ptr
  end subroutine

  !> Pushes changed knobs (albedo, emissivity, zenith angles, photon_scale_factor ...) to the device.
  subroutine radtran_cuda_sync_params(rc, params, err)
    type(RadtranCuda), intent(inout) :: rc
    type(c_params), intent(in) :: params
    character(:), allocatable, intent(out) :: err
    character(kind=c_char) :: err_c(err_len+1)
    call clima_rt_set_params(rc%handle, params, err_c)
    call copy_err(err_c, err)
  end subroutine

  !> Batched `Radtran%radiate` + `Radtran%TOA_fluxes` (clima_radtran.f90:221-342): the arguments
  !> are those of `radiate` with a trailing column dimension.
  subroutine radtran_radiate_batch(rc, T_surface, T, P, densities, dz, pdensities, radii, &
                                   compute_solar, compute_opacity, res, err)
    type(RadtranCuda), intent(inout) :: rc
    real(dp), intent(in) :: T_surface(:)         !! (ncol) K
    real(dp), intent(in) :: T(:,:), P(:,:)       !! (nz,ncol) K, bars
    real(dp), intent(in) :: densities(:,:,:)     !! (nz,ng,ncol) molecules/cm3
    real(dp), intent(in) :: dz(:,:)              !! (nz,ncol) cm
    real(dp), optional, target, intent(in) :: pdensities(:,:,:), radii(:,:,:) !! (nz,np,ncol)
    logical, optional, intent(in) :: compute_solar, compute_opacity
    type(RadtranBatchResult), target, intent(inout) :: res
    character(:), allocatable, intent(out) :: err

    integer(c_int32_t) :: ncol, nz, ng, np, csol, cop
    type(c_out) :: o
    type(c_ptr) :: pd, rd
    character(kind=c_char) :: err_c(err_len+1)

    ncol = int(size(T,2), c_int32_t); nz = int(size(T,1), c_int32_t)
    ng = int(size(densities,2), c_int32_t); np = int(rc%np, c_int32_t)
    if (size(T_surface) /= ncol .or. any(shape(P) /= shape(T)) .or. any(shape(dz) /= shape(T)) &
        .or. size(densities,1) /= nz .or. size(densities,3) /= ncol) then
      err = '"radiate_batch": input arrays have inconsistent dimensions.'
      return
    endif
    csol = 1; cop = 1
    if (present(compute_solar)) csol = merge(1_c_int32_t, 0_c_int32_t, compute_solar)
    if (present(compute_opacity)) cop = merge(1_c_int32_t, 0_c_int32_t, compute_opacity)
    pd = c_null_ptr; rd = c_null_ptr
    if (present(pdensities)) pd = c_loc(pdensities)
    if (present(radii)) rd = c_loc(radii)

    if (allocated(res%ISR)) then
      if (size(res%ISR) /= ncol .or. size(res%f_total,1) /= nz+1) then
        deallocate(res%fup_n_ir, res%fdn_n_ir, res%fup_n_sol, res%fdn_n_sol, res%f_total)
        deallocate(res%ISR, res%OLR, res%status)
      endif
    endif
    if (.not. allocated(res%ISR)) then
      allocate(res%fup_n_ir(nz+1,ncol), res%fdn_n_ir(nz+1,ncol))
      allocate(res%fup_n_sol(nz+1,ncol), res%fdn_n_sol(nz+1,ncol), res%f_total(nz+1,ncol))
      allocate(res%ISR(ncol), res%OLR(ncol), res%status(ncol))
    endif
    o%fup_n_ir = c_loc(res%fup_n_ir); o%fdn_n_ir = c_loc(res%fdn_n_ir)
    o%fup_n_sol = c_loc(res%fup_n_sol); o%fdn_n_sol = c_loc(res%fdn_n_sol)
    o%f_total = c_loc(res%f_total); o%ISR = c_loc(res%ISR); o%OLR = c_loc(res%OLR)
    o%fup_a_ir = c_null_ptr; o%fdn_a_ir = c_null_ptr; o%tau_band_ir = c_null_ptr
    o%fup_a_sol = c_null_ptr; o%fdn_a_sol = c_null_ptr; o%amean_sol = c_null_ptr
    o%tau_band_sol = c_null_ptr
    o%status = c_loc(res%status)

    call clima_rt_radiate_batch(rc%handle, ncol, nz, ng, np, T_surface, T, P, densities, dz, &
                                pd, rd, csol, cop, o, err_c)
    call copy_err(err_c, err)
  end subroutine

  !> The perturbed-temperature radiative transfers of `AdiabatClimate_jacobian_from_base`
  !> (src/adiabat/clima_adiabat_solve.f90:798-820) in one call.  `T_in(:,c)` is the model-grid vector
  !> [T_surf, T(1:nz)] of perturbation `c` (the reference's `T_perturb`); `f_total(:,c)` receives
  !> `rad%f_total(2*i-1)`, i = 1..nz+1 (:731-733), without `surface_heat_flow`.  The handle must hold
  !> the base column (the objective evaluation that precedes the loop, :757).
  subroutine radtran_jacobian_batch(rc, T_in, f_total, err)
    type(RadtranCuda), intent(inout) :: rc
    real(dp), intent(in) :: T_in(:,:)        !! (nz+1,ncol) K
    real(dp), intent(out) :: f_total(:,:)    !! (nz+1,ncol) mW/m^2
    character(:), allocatable, intent(out) :: err

    integer(c_int32_t) :: ncol, nz
    real(dp), allocatable :: T_surface(:), T(:,:)
    character(kind=c_char) :: err_c(err_len+1)

    ncol = int(size(T_in,2), c_int32_t); nz = int(size(T_in,1) - 1, c_int32_t)
    if (any(shape(f_total) /= shape(T_in))) then
      err = '"jacobian_batch": f_total must have the shape of T_in.'
      return
    endif
    allocate(T_surface(ncol), T(nz,ncol))
    T_surface(:) = T_in(1,:)
    T(:,:) = T_in(2:,:)
    call clima_rt_jacobian_batch(rc%handle, ncol, nz, T_surface, T, f_total, err_c)
    call copy_err(err_c, err)
  end subroutine

end module
