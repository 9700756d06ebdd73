!> clima_radtran_pack -- `Radtran_pack_tables`: gathers the opacity tables held by a Fortran `Radtran`
!> object (`self%op`, type OpticalProperties, src/radtran/clima_radtran_types.f90:96-141) into the
!> flat, column-major arrays that `struct clima_rt_tables` (include/clima_rt.h) points at, and creates
!> the CUDA handle from them.  It is the Fortran-side half of row (b) of SURVEY.md section 8: with it,
!> `create_Radtran_2` (src/radtran/clima_radtran.f90:128-219) can end with
!>
!>     call Radtran_pack_tables(rad, rad%cuda_pack, rad%cuda, err)
!>
!> and `rad%radiate_batch` works on the very tables `rad%radiate` uses.
!>
!> UNCOMPILED SOURCE: neither the build container nor the GPU box has a Fortran compiler
!> (SURVEY.md F2).  Written against Fortran 2008 and the reference at v0.7.4; the C struct member
!> order is the one test_abi.py checks for the ctypes mirror.
!>
!> One change to the reference is needed: the interpolator classes keep their node values private
!> (src/dependencies/linear_interpolation_module.F90:54-76: `f`, `x`, `y` are private components).
!> The packer reads them through two accessors to be added to that module (4 lines each):
!>
!>     procedure, public :: nodes => nodes_1d      ! linear_interp_1d:  subroutine nodes_1d(me, f)
!>     procedure, public :: nodes => nodes_2d      ! linear_interp_2d:  subroutine nodes_2d(me, f)
!>       real(wp), intent(out) :: f(:) / f(:,:);   f = me%f
!>
!> (Evaluating the interpolator AT its nodes returns the same numbers for finite tables, but
!> 0 * (-inf) = NaN where a k-table holds log10(0); the accessor has no such corner.)
module clima_radtran_pack
  use iso_c_binding
  use clima_const, only: dp
  use clima_radtran, only: Radtran
  use clima_radtran_types, only: OpticalProperties, Ktable, Xsection, ParticleXsection
  use clima_radtran_cuda, only: c_tables, c_params, RadtranCuda, radtran_cuda_create, pack_params
  implicit none
  private

  public :: RadtranTablePack, Radtran_pack_tables

  !> One ragged list of arrays as C sees it: `const double *const *`.
  type :: PtrList
    type(c_ptr), allocatable :: p(:)
  end type

  !> Owns every flat copy `c_tables` points into; must outlive `clima_rt_create` only (the library
  !> uploads the tables to the device and keeps no host pointer), kept in the Radtran object anyway.
  type :: RadtranTablePack
    real(c_double), allocatable :: wavl(:), weights(:), ray_xs(:,:), photons_sol(:)
    integer(c_int32_t), allocatable :: k_sp_ind(:), k_npress(:), k_ntemp(:)
    integer(c_int32_t), allocatable :: cia_sp_ind(:,:), cia_dim(:), cia_ntemp(:)
    integer(c_int32_t), allocatable :: ray_sp_ind(:)
    integer(c_int32_t), allocatable :: pxs_sp_ind(:), pxs_dim(:), pxs_ntemp(:)
    integer(c_int32_t), allocatable :: part_p_ind(:), part_nrad(:)
    real(c_double), allocatable :: cont_temp(:), cont_H2O(:,:), cont_foreign(:,:)
    ! ragged members: one flat buffer + offsets, and the pointer lists handed to C
    real(c_double), allocatable :: buf(:)
    type(PtrList) :: k_log10P, k_temp, k_log10k, cia_temp, cia_data, pxs_temp, pxs_data
    type(PtrList) :: part_radii, part_w0, part_qext, part_gt
  end type

contains

  !> Fills `pk` from `rad%op`, `rad%ir`, `rad%sol`, builds `c_tables` and creates the CUDA handle.
  subroutine Radtran_pack_tables(rad, pk, rc, err, device)
    type(Radtran), target, intent(in) :: rad
    type(RadtranTablePack), target, intent(inout) :: pk
    type(RadtranCuda), intent(inout) :: rc
    character(:), allocatable, intent(out) :: err
    integer, optional, intent(in) :: device

    type(c_tables) :: t
    type(c_params) :: p
    integer :: i, l, g, nw, ngauss, n, pos, need
    real(dp), allocatable :: f1(:), f2(:,:)
    integer :: dev

    associate (op => rad%op)
    nw = op%nw
    if (op%nk < 1) then
      err = 'Radtran_pack_tables: the CUDA path needs at least one k-distribution'
      return
    endif
    ngauss = op%k(1)%ngauss

    ! ---- size of the ragged buffer
    need = 0
    do i = 1,op%nk
      need = need + op%k(i)%npress + op%k(i)%ntemp + ngauss*op%k(i)%npress*op%k(i)%ntemp*nw
    enddo
    do i = 1,op%ncia
      need = need + xs_size(op%cia(i), nw)
    enddo
    do i = 1,op%npxs
      need = need + xs_size(op%pxs(i), nw)
    enddo
    do i = 1,op%npart
      need = need + op%part(i)%nrad*(1 + 3*nw)
    enddo
    if (allocated(pk%buf)) deallocate(pk%buf)
    allocate(pk%buf(max(need,1)))
    pos = 1

    ! ---- grid, Gauss weights, channels
    pk%wavl = op%wavl                       ! (nw+1) nm
    pk%weights = op%k(1)%weights            ! all tables share them (types_create.f90:379-389)
    t%nw = nw; t%wavl = c_loc(pk%wavl)
    t%ng = rad%ng; t%np = rad%np
    t%nk = op%nk; t%ngauss = ngauss; t%weights = c_loc(pk%weights)

    ! ---- k-distributions: log10k(g,l)%f(npress,ntemp) -> (ngauss,npress,ntemp,nw), g fastest
    !      (the layout of the HDF5 file, types_create.f90:1291-1358, and of the device table)
    allocate(pk%k_sp_ind(op%nk), pk%k_npress(op%nk), pk%k_ntemp(op%nk))
    allocate(pk%k_log10P%p(op%nk), pk%k_temp%p(op%nk), pk%k_log10k%p(op%nk))
    do i = 1,op%nk
      associate (k => op%k(i))
      if (k%ngauss /= ngauss) then
        err = 'Radtran_pack_tables: k-tables with different numbers of Gauss points'
        return
      endif
      pk%k_sp_ind(i) = k%sp_ind; pk%k_npress(i) = k%npress; pk%k_ntemp(i) = k%ntemp
      pk%k_log10P%p(i) = put(k%log10P)
      pk%k_temp%p(i) = put(k%temp)
      pk%k_log10k%p(i) = c_loc(pk%buf(pos))
      allocate(f2(k%npress,k%ntemp))
      do l = 1,nw
        do g = 1,ngauss
          call k%log10k(g,l)%nodes(f2)
          ! element (g, ip, it, l) of a column-major (ngauss,npress,ntemp,nw) array
          do n = 1,k%ntemp
            pk%buf(pos + (g-1) + ngauss*((n-1)*k%npress + (l-1)*k%npress*k%ntemp) : &
                   pos + (g-1) + ngauss*((n-1)*k%npress + (l-1)*k%npress*k%ntemp + k%npress - 1) : ngauss) = f2(:,n)
          enddo
        enddo
      enddo
      deallocate(f2)
      pos = pos + ngauss*k%npress*k%ntemp*nw
      end associate
    enddo
    t%k_sp_ind = c_loc(pk%k_sp_ind); t%k_npress = c_loc(pk%k_npress); t%k_ntemp = c_loc(pk%k_ntemp)
    t%k_log10P = c_loc(pk%k_log10P%p); t%k_temp = c_loc(pk%k_temp%p); t%k_log10k = c_loc(pk%k_log10k%p)

    ! ---- CIA (two species each) and photolysis cross sections (one species each)
    t%ncia = op%ncia
    allocate(pk%cia_sp_ind(2,max(op%ncia,1)), pk%cia_dim(max(op%ncia,1)), pk%cia_ntemp(max(op%ncia,1)))
    allocate(pk%cia_temp%p(max(op%ncia,1)), pk%cia_data%p(max(op%ncia,1)))
    do i = 1,op%ncia
      pk%cia_sp_ind(:,i) = op%cia(i)%sp_ind(1:2)
      call put_xs(op%cia(i), pk%cia_dim(i), pk%cia_ntemp(i), pk%cia_temp%p(i), pk%cia_data%p(i))
    enddo
    t%cia_sp_ind = c_loc(pk%cia_sp_ind); t%cia_dim = c_loc(pk%cia_dim); t%cia_ntemp = c_loc(pk%cia_ntemp)
    t%cia_temp = c_loc(pk%cia_temp%p); t%cia_data = c_loc(pk%cia_data%p)

    t%npxs = op%npxs
    allocate(pk%pxs_sp_ind(max(op%npxs,1)), pk%pxs_dim(max(op%npxs,1)), pk%pxs_ntemp(max(op%npxs,1)))
    allocate(pk%pxs_temp%p(max(op%npxs,1)), pk%pxs_data%p(max(op%npxs,1)))
    do i = 1,op%npxs
      pk%pxs_sp_ind(i) = op%pxs(i)%sp_ind(1)
      call put_xs(op%pxs(i), pk%pxs_dim(i), pk%pxs_ntemp(i), pk%pxs_temp%p(i), pk%pxs_data%p(i))
    enddo
    t%pxs_sp_ind = c_loc(pk%pxs_sp_ind); t%pxs_dim = c_loc(pk%pxs_dim); t%pxs_ntemp = c_loc(pk%pxs_ntemp)
    t%pxs_temp = c_loc(pk%pxs_temp%p); t%pxs_data = c_loc(pk%pxs_data%p)

    ! ---- Rayleigh: xs_0d per species (types_create.f90:1048-1088) -> (nw,nray)
    t%nray = op%nray
    allocate(pk%ray_sp_ind(max(op%nray,1)), pk%ray_xs(nw,max(op%nray,1)))
    do i = 1,op%nray
      pk%ray_sp_ind(i) = op%ray(i)%sp_ind(1)
      pk%ray_xs(:,i) = op%ray(i)%xs_0d
    enddo
    t%ray_sp_ind = c_loc(pk%ray_sp_ind); t%ray_xs = c_loc(pk%ray_xs)

    ! ---- water continuum (clima_radtran_types.f90:68-77)
    t%has_cont = 0; t%cont_LH2O = 0; t%cont_ntemp = 0
    t%cont_temp = c_null_ptr; t%cont_log10_H2O = c_null_ptr; t%cont_log10_foreign = c_null_ptr
    if (allocated(op%cont)) then
      t%has_cont = 1; t%cont_LH2O = op%cont%LH2O; t%cont_ntemp = op%cont%ntemp
      pk%cont_temp = op%cont%temp
      allocate(pk%cont_H2O(op%cont%ntemp,nw), pk%cont_foreign(op%cont%ntemp,nw), f1(op%cont%ntemp))
      do l = 1,nw
        call op%cont%log10_xs_H2O(l)%nodes(f1);     pk%cont_H2O(:,l) = f1
        call op%cont%log10_xs_foreign(l)%nodes(f1); pk%cont_foreign(:,l) = f1
      enddo
      deallocate(f1)
      t%cont_temp = c_loc(pk%cont_temp)
      t%cont_log10_H2O = c_loc(pk%cont_H2O); t%cont_log10_foreign = c_loc(pk%cont_foreign)
    endif

    ! ---- particles (clima_radtran_types.f90:57-66): w0, qext, gt (nrad,nw)
    t%npart = op%npart
    allocate(pk%part_p_ind(max(op%npart,1)), pk%part_nrad(max(op%npart,1)))
    allocate(pk%part_radii%p(max(op%npart,1)), pk%part_w0%p(max(op%npart,1)))
    allocate(pk%part_qext%p(max(op%npart,1)), pk%part_gt%p(max(op%npart,1)))
    do i = 1,op%npart
      associate (pt => op%part(i))
      pk%part_p_ind(i) = pt%p_ind; pk%part_nrad(i) = pt%nrad
      pk%part_radii%p(i) = put(pt%radii)
      allocate(f1(pt%nrad))
      pk%part_w0%p(i) = c_loc(pk%buf(pos))
      do l = 1,nw
        call pt%w0(l)%nodes(f1); pk%buf(pos:pos+pt%nrad-1) = f1; pos = pos + pt%nrad
      enddo
      pk%part_qext%p(i) = c_loc(pk%buf(pos))
      do l = 1,nw
        call pt%qext(l)%nodes(f1); pk%buf(pos:pos+pt%nrad-1) = f1; pos = pos + pt%nrad
      enddo
      pk%part_gt%p(i) = c_loc(pk%buf(pos))
      do l = 1,nw
        call pt%gt(l)%nodes(f1); pk%buf(pos:pos+pt%nrad-1) = f1; pos = pos + pt%nrad
      enddo
      deallocate(f1)
      end associate
    enddo
    t%part_p_ind = c_loc(pk%part_p_ind); t%part_nrad = c_loc(pk%part_nrad)
    t%part_radii = c_loc(pk%part_radii%p); t%part_w0 = c_loc(pk%part_w0%p)
    t%part_qext = c_loc(pk%part_qext%p); t%part_gt = c_loc(pk%part_gt%p)

    ! ---- channels (types_create.f90:250-268) and the stellar photon flux (clima_radtran.f90:63-65)
    t%ir_ind_start = rad%ir%ind_start; t%ir_ind_end = rad%ir%ind_end
    t%sol_ind_start = rad%sol%ind_start; t%sol_ind_end = rad%sol%ind_end
    pk%photons_sol = rad%photons_sol
    t%photons_sol = c_loc(pk%photons_sol)
    end associate

    ! ---- run-time knobs, then the handle
    ! (the vectors live in `rad`, which is a target: pack_params takes their addresses)
    call pack_params(rad%diurnal_fac, rad%photon_scale_factor, rad%ir_tau_min, rad%has_hard_surface, &
                     rad%zenith_u, rad%zenith_weights, rad%surface_albedo, rad%surface_emissivity, p)
    dev = 0
    if (present(device)) dev = device
    call radtran_cuda_create(rc, t, p, dev, err)

  contains

    !> number of doubles one cross section needs in the ragged buffer
    pure function xs_size(xs, nw) result(n)
      type(Xsection), intent(in) :: xs
      integer, intent(in) :: nw
      integer :: n
      if (xs%dim == 0) then
        n = 1 + nw
      else
        n = xs%ntemp*(1 + nw)
      endif
    end function

    !> copies `a` into the ragged buffer and returns its address
    function put(a) result(ptr)
      real(dp), intent(in) :: a(:)
      type(c_ptr) :: ptr
      pk%buf(pos:pos+size(a)-1) = a
      ptr = c_loc(pk%buf(pos))
      pos = pos + size(a)
    end function

    !> dim 0: xs_0d(nw); dim 1: log10 xs (ntemp,nw) read out of the nw interpolators
    subroutine put_xs(xs, xdim, xnt, ptemp, pdata)
      type(Xsection), intent(in) :: xs
      integer(c_int32_t), intent(out) :: xdim, xnt
      type(c_ptr), intent(out) :: ptemp, pdata
      real(dp), allocatable :: f(:)
      integer :: ll
      xdim = xs%dim
      if (xs%dim == 0) then
        xnt = 0
        ptemp = put([0.0_dp])
        pdata = put(xs%xs_0d)
      else
        xnt = xs%ntemp
        ptemp = put(xs%temp)
        pdata = c_loc(pk%buf(pos))
        allocate(f(xs%ntemp))
        do ll = 1,size(xs%log10_xs_1d)
          call xs%log10_xs_1d(ll)%nodes(f)
          pk%buf(pos:pos+xs%ntemp-1) = f
          pos = pos + xs%ntemp
        enddo
      endif
    end subroutine

  end subroutine

end module
