!> clima_batch_wrappers -- the `bind(c)` wrappers that expose the batched entry points to the Cython
!> layer, in the style of clima/fortran/Radtran.f90:77-107 and clima/fortran/AdiabatClimate.f90:169-190
!> (opaque pointer by value, scalars by reference, dims before arrays, `err(err_len+1)` last).
!> To be `include`d in clima/fortran/clima_c_api.f90 next to Radtran.f90 / AdiabatClimate.f90.
!>
!> UNCOMPILED SOURCE (no Fortran compiler on either box, SURVEY.md F2).
!>
!> radtran_radiate_batch_wrapper        Radtran%radiate_batch  (INTEGRATION.md section 2c)
!> adiabatclimate_toa_fluxes_batch_wrapper
!>     `AdiabatClimate%TOA_fluxes` for ncol (T_surf, P_i_surf) pairs: `make_profile` +
!>     `copy_atm_to_radiative_grid` per column on the host (unchanged reference code,
!>     src/adiabat/clima_adiabat.f90:401-472,729-773), then ONE `rad%radiate_batch` for all columns
!>     instead of ncol calls of `rad%TOA_fluxes` (:474-520).  This is row N3 of SURVEY.md section 8(f);
!>     batching `make_profile` itself is row N4 and is not part of this repository.

subroutine radtran_radiate_batch_wrapper(ptr, ncol, nz, ng, T_surface, T, P, densities, dz, &
                                         compute_solar, fup_n_ir, fdn_n_ir, fup_n_sol, fdn_n_sol, &
                                         f_total, ISR, OLR, err) bind(c)
  use clima, only: Radtran
  use clima_radtran_cuda, only: RadtranBatchResult
  type(c_ptr), value, intent(in) :: ptr
  integer(c_int), intent(in) :: ncol, nz, ng
  real(c_double), intent(in) :: T_surface(ncol), T(nz,ncol), P(nz,ncol), densities(nz,ng,ncol), dz(nz,ncol)
  logical(c_bool), intent(in) :: compute_solar
  real(c_double), intent(out) :: fup_n_ir(nz+1,ncol), fdn_n_ir(nz+1,ncol)
  real(c_double), intent(out) :: fup_n_sol(nz+1,ncol), fdn_n_sol(nz+1,ncol), f_total(nz+1,ncol)
  real(c_double), intent(out) :: ISR(ncol), OLR(ncol)
  character(c_char), intent(out) :: err(err_len+1)

  character(:), allocatable :: err_f
  type(Radtran), pointer :: rad
  type(RadtranBatchResult) :: res
  logical :: compute_solar_f

  call c_f_pointer(ptr, rad)
  compute_solar_f = compute_solar
  call rad%radiate_batch(T_surface, T, P, densities, dz, compute_solar=compute_solar_f, res=res, err=err_f)

  err(1) = c_null_char
  if (allocated(err_f)) then
    call copy_string_ftoc(err_f, err)
    return
  endif
  fup_n_ir = res%fup_n_ir; fdn_n_ir = res%fdn_n_ir
  fup_n_sol = res%fup_n_sol; fdn_n_sol = res%fdn_n_sol
  f_total = res%f_total; ISR = res%ISR; OLR = res%OLR
end subroutine

subroutine adiabatclimate_toa_fluxes_batch_wrapper(ptr, ncol, T_surf, ng, P_i_surf, ISR, OLR, err) bind(c)
  use clima, only: AdiabatClimate
  use clima_radtran_cuda, only: RadtranBatchResult
  type(c_ptr), value, intent(in) :: ptr
  integer(c_int), intent(in) :: ncol
  real(c_double), intent(in) :: T_surf(ncol)
  integer(c_int), intent(in) :: ng
  real(c_double), intent(in) :: P_i_surf(ng,ncol)
  real(c_double), intent(out) :: ISR(ncol), OLR(ncol)
  character(c_char), intent(out) :: err(err_len+1)

  character(:), allocatable :: err_f
  type(AdiabatClimate), pointer :: c
  type(RadtranBatchResult) :: res
  real(c_double), allocatable :: T(:,:), P(:,:), dens(:,:,:), dz(:,:)
  integer :: i, nzr

  call c_f_pointer(ptr, c)
  err(1) = c_null_char
  nzr = c%rad%nz                      ! the radiative grid of copy_atm_to_radiative_grid: 2*nz + 2
  allocate(T(nzr,ncol), P(nzr,ncol), dens(nzr,c%sp%ng,ncol), dz(nzr,ncol))
  do i = 1,ncol
    ! column construction stays the reference's (row N4): pseudo-adiabat, then the doubled grid
    call c%make_profile(T_surf(i), P_i_surf(:,i), err_f)
    if (allocated(err_f)) then
      call copy_string_ftoc(err_f, err)
      return
    endif
    call c%copy_atm_to_radiative_grid()
    T(:,i) = c%T_r; P(:,i) = c%P_r/1.0e6_dp; dens(:,:,i) = c%densities_r; dz(:,i) = c%dz_r
  enddo
  call c%rad%radiate_batch(T_surf, T, P, dens, dz, res=res, err=err_f)
  if (allocated(err_f)) then
    call copy_string_ftoc(err_f, err)
    return
  endif
  ISR = res%ISR; OLR = res%OLR
end subroutine
