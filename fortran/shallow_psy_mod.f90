!> The application ("PSy") layer of a GOcean shallow-water model over device-resident fd_lib
!! fields: what replaces the `do jj / do ji / call kernel_code(ji,jj,fld%data,...)` nests of
!! SURVEY CS-3 (src/kernel_mod.f90:35-37, src/infrastructure_mod.f90:32-41).
!! Two forms: one launch per reference-style kernel (same sweep bounds as the loops), or the
!! whole step fused into one sweep (96 B/point instead of 368).
!! SOURCE ONLY -- never compiled here (no Fortran compiler on either box).
module shallow_psy_mod
  use iso_c_binding
  use kind_params_mod, only: go_wp
  use field_mod, only: r2d_field, copy_field
  use parallel_utils_mod, only: get_num_ranks
  use field_device_mod
  use dlesm_b200_mod
  implicit none
  private
  public :: sw_step_fused, sw_step2_fused, sw_step_unfused

contains

  !> One model time step, fused.  On return unew,vnew,pnew hold step n+1 with valid halos and
  !! uold,vold,pold the filtered step n; the caller rotates (u,unew) etc. by swapping the
  !! r2d_field objects -- no copy_field.
  subroutine sw_step_fused(u, v, p, uold, vold, pold, unew, vnew, pnew, tdt, alpha, first)
    type(r2d_field), intent(inout) :: u, v, p, uold, vold, pold, unew, vnew, pnew
    real(go_wp), intent(in) :: tdt, alpha
    logical, intent(in) :: first
    integer(c_int) :: ifirst
    ifirst = 0
    if (first) ifirst = 1
    ! sweep bounds: the T-point internal region (== every stagger's, GO_OFFSET_SW periodic,
    ! src/field_mod.f90:694-729,817-826,935-938,1040-1060)
    call dlesm_check(dlesm_sw_step_fused(u%device_ptr, v%device_ptr, p%device_ptr,            &
         uold%device_ptr, vold%device_ptr, pold%device_ptr,                                    &
         unew%device_ptr, vnew%device_ptr, pnew%device_ptr,                                    &
         real(p%grid%dx, c_double), real(p%grid%dy, c_double), real(tdt, c_double),            &
         real(alpha, c_double), ifirst,                                                        &
         int(p%internal%xstart, c_int), int(p%internal%xstop, c_int),                          &
         int(p%internal%ystart, c_int), int(p%internal%ystop, c_int),                          &
         p%grid%tmask_device, 0_c_int), 'sw_step_fused')
    call fill_halos_3(unew, vnew, pnew)
  end subroutine sw_step_fused

  !> TWO model time steps in one sweep (temporal blocking): the state crosses HBM once for both.
  !! Needs grid%decompose(..., halo_width=2); u,v,p with halos two cells deep and uold,vold,pold
  !! one cell deep on entry (fill_halos_deep below); never the first step (tdt = 2*dt).  On return
  !! u2,v2,p2 hold step n+2 and uold2,vold2,pold2 the twice-filtered old fields, halos two cells
  !! deep; the caller rotates (u,u2), (uold,uold2) etc. by swapping the r2d_field objects.
  subroutine sw_step2_fused(u, v, p, uold, vold, pold, u2, v2, p2, uold2, vold2, pold2, tdt, alpha)
    type(r2d_field), intent(inout) :: u, v, p, uold, vold, pold, u2, v2, p2, uold2, vold2, pold2
    real(go_wp), intent(in) :: tdt, alpha
    call dlesm_check(dlesm_sw_step2_fused(u%device_ptr, v%device_ptr, p%device_ptr,           &
         uold%device_ptr, vold%device_ptr, pold%device_ptr,                                    &
         u2%device_ptr, v2%device_ptr, p2%device_ptr,                                          &
         uold2%device_ptr, vold2%device_ptr, pold2%device_ptr,                                 &
         real(p%grid%dx, c_double), real(p%grid%dy, c_double), real(tdt, c_double),            &
         real(alpha, c_double),                                                                &
         int(p%internal%xstart, c_int), int(p%internal%xstop, c_int),                          &
         int(p%internal%ystart, c_int), int(p%internal%ystop, c_int), c_null_ptr, 0_c_int), 'sw_step2_fused')
    call fill_halos_deep(u2);     call fill_halos_deep(v2);     call fill_halos_deep(p2)
    call fill_halos_deep(uold2);  call fill_halos_deep(vold2);  call fill_halos_deep(pold2)
  end subroutine sw_step2_fused

  !> halos two cells deep: wrapped copies on one rank, halo_exchange(depth=2) on several
  !! (the DLESM_EXT_HALO_DEPTH extension; the reference ignores depth, src/field_mod.f90:1226-1229)
  subroutine fill_halos_deep(f)
    type(r2d_field), intent(inout) :: f
    if (get_num_ranks() == 1) then
       call dlesm_check(dlesm_dev_periodic_halos_deep(f%device_ptr,                            &
            int(f%internal%xstart, c_int), int(f%internal%xstop, c_int),                       &
            int(f%internal%ystart, c_int), int(f%internal%ystop, c_int), 2_c_int), 'halos_deep')
    else
       call f%halo_exchange(2)
    end if
  end subroutine fill_halos_deep

  !> The same step, one launch per kernel, each over its output field's internal region.
  subroutine sw_step_unfused(u, v, p, uold, vold, pold, unew, vnew, pnew, cu, cv, z, h, &
                             tdt, alpha, first)
    type(r2d_field), intent(inout) :: u, v, p, uold, vold, pold, unew, vnew, pnew, cu, cv, z, h
    real(go_wp), intent(in) :: tdt, alpha
    logical, intent(in) :: first
    type(c_ptr) :: tm
    real(c_double) :: dx, dy
    tm = p%grid%tmask_device
    dx = p%grid%dx;  dy = p%grid%dy
    call dlesm_check(dlesm_compute_cu(cu%device_ptr, p%device_ptr, u%device_ptr, b(cu,1), b(cu,2), b(cu,3), b(cu,4), tm), 'cu')
    call dlesm_check(dlesm_compute_cv(cv%device_ptr, p%device_ptr, v%device_ptr, b(cv,1), b(cv,2), b(cv,3), b(cv,4), tm), 'cv')
    call dlesm_check(dlesm_compute_z(z%device_ptr, p%device_ptr, u%device_ptr, v%device_ptr, dx, dy, &
                                     b(z,1), b(z,2), b(z,3), b(z,4), tm), 'z')
    call dlesm_check(dlesm_compute_h(h%device_ptr, p%device_ptr, u%device_ptr, v%device_ptr, &
                                     b(h,1), b(h,2), b(h,3), b(h,4), tm), 'h')
    call fill_halos_3(cu, cv, z);  call fill_halos_1(h)
    call dlesm_check(dlesm_compute_unew(unew%device_ptr, uold%device_ptr, z%device_ptr, cv%device_ptr, h%device_ptr, &
         real(tdt, c_double), dx, b(unew,1), b(unew,2), b(unew,3), b(unew,4), tm, u%device_ptr), 'unew')
    call dlesm_check(dlesm_compute_vnew(vnew%device_ptr, vold%device_ptr, z%device_ptr, cu%device_ptr, h%device_ptr, &
         real(tdt, c_double), dy, b(vnew,1), b(vnew,2), b(vnew,3), b(vnew,4), tm, v%device_ptr), 'vnew')
    call dlesm_check(dlesm_compute_pnew(pnew%device_ptr, pold%device_ptr, cu%device_ptr, cv%device_ptr, &
         real(tdt, c_double), dx, dy, b(pnew,1), b(pnew,2), b(pnew,3), b(pnew,4), tm, p%device_ptr), 'pnew')
    call fill_halos_3(unew, vnew, pnew)
    if (.not. first) then
       call dlesm_check(dlesm_time_smooth(u%device_ptr, unew%device_ptr, uold%device_ptr, real(alpha, c_double), &
            b(u,1), b(u,2), b(u,3), b(u,4), tm, 0_c_int), 'smooth u')     ! GO_U_POINTS
       call dlesm_check(dlesm_time_smooth(v%device_ptr, vnew%device_ptr, vold%device_ptr, real(alpha, c_double), &
            b(v,1), b(v,2), b(v,3), b(v,4), tm, 1_c_int), 'smooth v')     ! GO_V_POINTS
       call dlesm_check(dlesm_time_smooth(p%device_ptr, pnew%device_ptr, pold%device_ptr, real(alpha, c_double), &
            b(p,1), b(p,2), b(p,3), b(p,4), tm, 2_c_int), 'smooth p')     ! GO_T_POINTS
    end if
    call copy_field(unew, u);  call copy_field(vnew, v);  call copy_field(pnew, p)   ! device-resident: one D2D pass each
  end subroutine sw_step_unfused

  !> internal-region bound k of a field as c_int (1 xstart, 2 xstop, 3 ystart, 4 ystop)
  pure function b(f, k) result(v)
    type(r2d_field), intent(in) :: f
    integer, intent(in) :: k
    integer(c_int) :: v
    select case (k)
    case (1); v = int(f%internal%xstart, c_int)
    case (2); v = int(f%internal%xstop, c_int)
    case (3); v = int(f%internal%ystart, c_int)
    case default; v = int(f%internal%ystop, c_int)
    end select
  end function b

  subroutine fill_halos_3(f1, f2, f3)
    type(r2d_field), intent(inout) :: f1, f2, f3
    if (get_num_ranks() > 1) then
       call dev_halo_exchange_3(f1, f2, f3)
    else
       call dev_periodic_halos(f1);  call dev_periodic_halos(f2);  call dev_periodic_halos(f3)
    end if
  end subroutine fill_halos_3

  subroutine fill_halos_1(f)
    type(r2d_field), intent(inout) :: f
    if (get_num_ranks() > 1) then
       call f%halo_exchange(1)
    else
       call dev_periodic_halos(f)
    end if
  end subroutine fill_halos_1

end module shallow_psy_mod
