!> clima_adiabat_cuda -- row N4 of SURVEY.md section 8(f): the batched `AdiabatClimate%TOA_fluxes` with
!> `make_profile` itself on the device.  The reference-side binding of
!>   clima_rt_adiabat_set, clima_rt_make_profile_batch, clima_rt_adiabat_toa_fluxes_batch  (include/clima_rt.h)
!>
!> UNCOMPILED SOURCE (no Fortran compiler on either box, SURVEY.md F2), written against
!> src/adiabat/clima_adiabat.f90:13-135 (the members of AdiabatClimate), src/clima_types.f90:115-150
!> (Species, Gas, ThermodynamicData) and src/clima_saturationdata.f90:16-28 (SaturationData).
!>
!> `AdiabatClimate_TOA_fluxes_batch_device` is to be added to src/adiabat/clima_adiabat.f90 as a type-bound
!> procedure next to AdiabatClimate_TOA_fluxes (:776-799) and to the host-profile variant of
!> fortran/clima_batch_wrappers.f90; `self%rad%cuda` is the RadtranCuda member that
!> fortran/clima_radtran_cuda.f90 adds to `type Radtran`.
module clima_adiabat_cuda
  use iso_c_binding
  use clima_const, only: dp
  implicit none
  private
  public :: clima_rt_adiabat, clima_rt_profile_out, adiabat_to_device
  public :: clima_rt_adiabat_set, clima_rt_make_profile_batch, clima_rt_adiabat_toa_fluxes_batch

  integer, parameter :: max_gases = 16, max_temps = 4   ! CLIMA_RT_ADIABAT_MAX_GASES / _MAX_TEMPS

  type, bind(c) :: clima_rt_adiabat
    integer(c_int32_t) :: ng, nz, double_radiative_grid
    type(c_ptr) :: mass, thermo_model, thermo_ntemps, thermo_temps, thermo_data, has_sat, sat, RH
    real(c_double) :: planet_mass, planet_radius, P_top, T_trop, rtol, atol, reference_pressure
  end type

  type, bind(c) :: clima_rt_profile_out
    type(c_ptr) :: P, T, z, dz, f_i, f_i_surf, N_surface, P_surf, P_trop, status, nfcn
  end type

  interface
    subroutine clima_rt_adiabat_set(handle, adiabat, err) bind(c, name='clima_rt_adiabat_set')
      import :: c_ptr, c_char, clima_rt_adiabat
      type(c_ptr), value :: handle
      type(clima_rt_adiabat), intent(in) :: adiabat
      character(c_char), intent(out) :: err(*)
    end subroutine
    subroutine clima_rt_make_profile_batch(handle, ncol, T_surf, P_i_surf, profile, err) &
        bind(c, name='clima_rt_make_profile_batch')
      import :: c_ptr, c_char, c_int32_t, c_double, clima_rt_profile_out
      type(c_ptr), value :: handle
      integer(c_int32_t), intent(in) :: ncol
      real(c_double), intent(in) :: T_surf(*), P_i_surf(*)
      type(clima_rt_profile_out), intent(in) :: profile
      character(c_char), intent(out) :: err(*)
    end subroutine
    subroutine clima_rt_adiabat_toa_fluxes_batch(handle, ncol, T_surf, P_i_surf, profile, out, err) &
        bind(c, name='clima_rt_adiabat_toa_fluxes_batch')
      use clima_radtran_cuda, only: c_out
      import :: c_ptr, c_char, c_int32_t, c_double
      type(c_ptr), value :: handle
      integer(c_int32_t), intent(in) :: ncol
      real(c_double), intent(in) :: T_surf(*), P_i_surf(*)
      type(c_ptr), value :: profile        ! a clima_rt_profile_out, or c_null_ptr
      type(c_out), intent(in) :: out
      character(c_char), intent(out) :: err(*)
    end subroutine
  end interface

contains

  !> Packs what make_profile reads from the object and hands it to the device.  Call after any change of
  !> RH, P_top, T_trop, rtol, atol (the setters of clima/fortran/AdiabatClimate.f90).
  subroutine adiabat_to_device(handle, sp, nz, double_radiative_grid, planet_mass, planet_radius, &
                               P_top, T_trop, RH, rtol, atol, reference_pressure, err)
    use clima_types, only: Species
    type(c_ptr), intent(in) :: handle
    type(Species), intent(in) :: sp
    integer, intent(in) :: nz
    logical, intent(in) :: double_radiative_grid
    real(dp), intent(in) :: planet_mass, planet_radius, P_top, T_trop, RH(:), rtol, atol, reference_pressure
    character(:), allocatable, intent(out) :: err

    type(clima_rt_adiabat) :: a
    real(c_double), target :: mass(max_gases), temps(max_temps+1,max_gases), coef(9,max_temps,max_gases)
    real(c_double), target :: sat(11,max_gases), RH_c(max_gases)
    integer(c_int32_t), target :: model(max_gases), ntemps(max_gases), has_sat(max_gases)
    character(kind=c_char) :: err_c(1025)
    integer :: i, k, n

    if (sp%ng > max_gases) then
      err = 'adiabat_to_device: more than 16 gases'
      return
    endif
    temps = 0.0_dp; coef = 0.0_dp; sat = 0.0_dp; has_sat = 0
    do i = 1,sp%ng
      mass(i) = sp%g(i)%mass
      model(i) = sp%g(i)%thermo%dtype          ! ShomatePolynomial = 1, Nasa9Polynomial = 2 (clima_types.f90:111-112)
      ntemps(i) = sp%g(i)%thermo%ntemps
      if (ntemps(i) > max_temps) then
        err = 'adiabat_to_device: more than 4 temperature ranges for "'//sp%g(i)%name//'"'
        return
      endif
      temps(1:ntemps(i)+1,i) = sp%g(i)%thermo%temps
      n = size(sp%g(i)%thermo%data,1)
      do k = 1,ntemps(i)
        coef(1:n,k,i) = sp%g(i)%thermo%data(:,k)
      enddo
      if (allocated(sp%g(i)%sat)) then
        has_sat(i) = 1
        associate (s => sp%g(i)%sat)
        sat(:,i) = [s%mu, s%T_ref, s%P_ref, s%T_triple, s%T_critical, s%a_v, s%b_v, s%a_s, s%b_s, s%a_c, s%b_c]
        end associate
      endif
      RH_c(i) = RH(i)
    enddo
    a%ng = sp%ng; a%nz = nz; a%double_radiative_grid = merge(1, 0, double_radiative_grid)
    a%mass = c_loc(mass); a%thermo_model = c_loc(model); a%thermo_ntemps = c_loc(ntemps)
    a%thermo_temps = c_loc(temps); a%thermo_data = c_loc(coef); a%has_sat = c_loc(has_sat)
    a%sat = c_loc(sat); a%RH = c_loc(RH_c)
    a%planet_mass = planet_mass; a%planet_radius = planet_radius
    a%P_top = P_top; a%T_trop = T_trop; a%rtol = rtol; a%atol = atol; a%reference_pressure = reference_pressure
    call clima_rt_adiabat_set(handle, a, err_c)
    call err_from_c(err_c, err)
  end subroutine

  subroutine err_from_c(err_c, err)
    character(kind=c_char), intent(in) :: err_c(:)
    character(:), allocatable, intent(out) :: err
    integer :: n
    n = 0
    do while (n < size(err_c) - 1)
      if (err_c(n+1) == c_null_char) exit
      n = n + 1
    enddo
    if (n > 0) then
      allocate(character(n) :: err)
      err = transfer(err_c(1:n), err)
    endif
  end subroutine

end module

!> To be added to src/adiabat/clima_adiabat.f90 (`procedure :: TOA_fluxes_batch_device => ...`): the loop
!>   do i = 1,ncol;  call c%TOA_fluxes(T_surf(i), P_i_surf(:,i), ISR(i), OLR(i), err);  enddo
!> of a sweep (examples/EarlyMars.ipynb, cell 7) as ONE device call.  Not supported here, as in
!> include/clima_rt.h: ocean solubility functions (they only fill N_ocean, which this entry does not return; the
!> host-profile variant in fortran/clima_batch_wrappers.f90 keeps them) and a T_surf-dependent albedo.  Particle profiles: forward the arguments of
!> AdiabatClimate_set_particle_density_and_radii to clima_rt_adiabat_set_particles (same argument list + handle); objects
!> with particles are refused below until that forwarding is added.
subroutine AdiabatClimate_TOA_fluxes_batch_device(self, T_surf, P_i_surf, ISR, OLR, err)
  use iso_c_binding
  use clima_adiabat_cuda
  use clima_radtran_cuda, only: c_out
  class(AdiabatClimate), intent(inout) :: self
  real(dp), intent(in) :: T_surf(:)      !! (ncol) K
  real(dp), intent(in) :: P_i_surf(:,:)  !! (ng,ncol) dynes/cm^2
  real(dp), target, intent(out) :: ISR(:), OLR(:)  !! (ncol) mW/m^2
  character(:), allocatable, intent(out) :: err

  type(c_out) :: out
  real(dp), allocatable, target :: fup_ir(:,:), fdn_ir(:,:), fup_sol(:,:), fdn_sol(:,:), f_total(:,:)
  character(kind=c_char) :: err_c(1025)
  integer :: ncol, i

  ncol = size(T_surf)
  if (size(P_i_surf,1) /= self%sp%ng .or. size(P_i_surf,2) /= ncol .or. size(ISR) /= ncol .or. size(OLR) /= ncol) then
    err = '"T_surf" and "P_i_surf" have inconsistent dimensions'
    return
  endif
  if (associated(self%albedo_fcn) .or. self%sp%np > 0) then
    err = 'TOA_fluxes_batch_device: albedo_fcn and particles need the host-profile variant'
    return
  endif
  do i = 1,self%sp%ng
    if (associated(self%ocean_fcns(i)%fcn)) then
      err = 'TOA_fluxes_batch_device: ocean solubility functions need the host-profile variant'
      return
    endif
  enddo
  call adiabat_to_device(self%rad%cuda%handle, self%sp, self%nz, self%double_radiative_grid, self%planet_mass, &
                         self%planet_radius, self%P_top, self%T_trop, self%RH, self%rtol, self%atol, &
                         self%reference_pressure, err)
  if (allocated(err)) return
  allocate(fup_ir(self%nz_r+1,ncol), fdn_ir(self%nz_r+1,ncol), fup_sol(self%nz_r+1,ncol), fdn_sol(self%nz_r+1,ncol))
  allocate(f_total(self%nz_r+1,ncol))
  out%fup_n_ir = c_loc(fup_ir); out%fdn_n_ir = c_loc(fdn_ir); out%fup_n_sol = c_loc(fup_sol)
  out%fdn_n_sol = c_loc(fdn_sol); out%f_total = c_loc(f_total); out%ISR = c_loc(ISR); out%OLR = c_loc(OLR)
  out%fup_a_ir = c_null_ptr; out%fdn_a_ir = c_null_ptr; out%tau_band_ir = c_null_ptr
  out%fup_a_sol = c_null_ptr; out%fdn_a_sol = c_null_ptr; out%amean_sol = c_null_ptr
  out%tau_band_sol = c_null_ptr; out%status = c_null_ptr
  call clima_rt_adiabat_toa_fluxes_batch(self%rad%cuda%handle, int(ncol, c_int32_t), T_surf, P_i_surf, c_null_ptr, &
                                         out, err_c)
  if (err_c(1) /= c_null_char) then
    block
      integer :: n
      n = 0
      do while (n < 1024)
        if (err_c(n+1) == c_null_char) exit
        n = n + 1
      enddo
      allocate(character(n) :: err)
      err = transfer(err_c(1:n), err)
    end block
  endif
end subroutine
