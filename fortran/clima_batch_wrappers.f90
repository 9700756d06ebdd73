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
!>     `make_profile` itself on the device (row N4) is fortran/clima_adiabat_cuda.f90.

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

!> To be added to src/adiabat/clima_adiabat.f90 as a type-bound procedure of AdiabatClimate
!> (`procedure :: TOA_fluxes_batch => AdiabatClimate_TOA_fluxes_batch`), next to
!> AdiabatClimate_TOA_fluxes (:776-799), whose steps it repeats per column up to the RT call:
!> `copy_atm_to_radiative_grid` is a private binding (:219), so this cannot live outside the module.
subroutine AdiabatClimate_TOA_fluxes_batch(self, T_surf, P_i_surf, ISR, OLR, err)
  use clima_radtran_cuda, only: RadtranBatchResult
  class(AdiabatClimate), intent(inout) :: self
  real(dp), intent(in) :: T_surf(:)      !! (ncol) K
  real(dp), intent(in) :: P_i_surf(:,:)  !! (ng,ncol) dynes/cm^2
  real(dp), intent(out) :: ISR(:), OLR(:)  !! (ncol) mW/m^2
  character(:), allocatable, intent(out) :: err

  type(RadtranBatchResult) :: res
  real(dp), allocatable :: T(:,:), P(:,:), dens(:,:,:), dz(:,:), pdens(:,:,:), radii(:,:,:)
  integer :: i, ncol

  ncol = size(T_surf)
  if (size(P_i_surf,2) /= ncol .or. size(ISR) /= ncol .or. size(OLR) /= ncol) then
    err = 'TOA_fluxes_batch: inconsistent number of columns'
    return
  endif
  if (associated(self%albedo_fcn)) then
    ! one set of run-time knobs per batched call: a surface albedo that depends on T_surf (:789-791)
    ! needs one call per albedo value
    err = 'TOA_fluxes_batch: "albedo_fcn" is set; call TOA_fluxes per column or batch columns of equal albedo'
    return
  endif
  allocate(T(self%nz_r,ncol), P(self%nz_r,ncol), dens(self%nz_r,self%sp%ng,ncol), dz(self%nz_r,ncol))
  allocate(pdens(self%nz_r,self%sp%np,ncol), radii(self%nz_r,self%sp%np,ncol))
  do i = 1,ncol
    ! column construction stays the reference's (row N4): pseudo-adiabat, then the doubled grid
    call self%make_profile(T_surf(i), P_i_surf(:,i), err)
    if (allocated(err)) return
    call self%copy_atm_to_radiative_grid()
    T(:,i) = self%T_r; P(:,i) = self%P_r/1.0e6_dp; dens(:,:,i) = self%densities_r; dz(:,i) = self%dz_r
    pdens(:,:,i) = self%pdensities_r; radii(:,:,i) = self%pradii_r
  enddo
  call self%rad%radiate_batch(T_surf, T, P, dens, dz, pdens, radii, res=res, err=err)
  if (allocated(err)) return
  ISR = res%ISR; OLR = res%OLR
end subroutine

subroutine adiabatclimate_toa_fluxes_batch_wrapper(ptr, ncol, T_surf, ng, P_i_surf, ISR, OLR, err) bind(c)
  use clima, only: AdiabatClimate
  type(c_ptr), value, intent(in) :: ptr
  integer(c_int), intent(in) :: ncol
  real(c_double), intent(in) :: T_surf(ncol)
  integer(c_int), intent(in) :: ng
  real(c_double), intent(in) :: P_i_surf(ng,ncol)
  real(c_double), intent(out) :: ISR(ncol), OLR(ncol)
  character(c_char), intent(out) :: err(err_len+1)

  character(:), allocatable :: err_f
  type(AdiabatClimate), pointer :: c

  call c_f_pointer(ptr, c)
  call c%TOA_fluxes_batch(T_surf, P_i_surf, ISR, OLR, err_f)
  err(1) = c_null_char
  if (allocated(err_f)) then
    call copy_string_ftoc(err_f, err)
  endif
end subroutine

!> Row N4: same signature, `make_profile` on the device (fortran/clima_adiabat_cuda.f90).
subroutine adiabatclimate_toa_fluxes_batch_device_wrapper(ptr, ncol, T_surf, ng, P_i_surf, ISR, OLR, err) bind(c)
  use clima, only: AdiabatClimate
  type(c_ptr), value, intent(in) :: ptr
  integer(c_int), intent(in) :: ncol
  real(c_double), intent(in) :: T_surf(ncol)
  integer(c_int), intent(in) :: ng
  real(c_double), intent(in) :: P_i_surf(ng,ncol)
  real(c_double), intent(out) :: ISR(ncol), OLR(ncol)
  character(c_char), intent(out) :: err(err_len+1)

  character(:), allocatable :: err_f
  type(AdiabatClimate), pointer :: c

  call c_f_pointer(ptr, c)
  call c%TOA_fluxes_batch_device(T_surf, P_i_surf, ISR, OLR, err_f)
  err(1) = c_null_char
  if (allocated(err_f)) then
    call copy_string_ftoc(err_f, err)
  endif
end subroutine
