!> Device residency for r2d_field through the reference's own seam
!! (data_on_device / device_ptr / read_from_device_c / write_to_device_c,
!! src/field_mod.f90:128-134,153): what a B200 build of field_mod adds.
!!
!! Either keep this as a separate module an application `use`s next to field_mod, or fold
!! the bodies into field_mod behind `if (fld%data_on_device)` (INTEGRATION.md shows where).
!! SOURCE ONLY -- no Fortran compiler on the build or GPU box; never compiled here.
module field_device_mod
  use iso_c_binding
  use kind_params_mod, only: go_wp
  use field_mod, only: r2d_field
  use region_mod, only: region_type
  use grid_mod, only: grid_type, GO_BC_PERIODIC
  use dlesm_b200_mod
  implicit none
  private
  public :: field_to_device, field_free_device
  public :: dev_copy_field, dev_copy_field_patch, dev_set_field, dev_field_checksum
  public :: dev_halo_exchange, dev_halo_exchange_3, dev_periodic_halos

contains

  !> Give `fld` device storage, install the two C hooks and upload the host values.
  !! After this fld%get_data() / read_from_device() / write_to_device() of the reference
  !! (src/field_mod.f90:407-559) work unchanged: they call the hooks installed here.
  subroutine field_to_device(fld)
    type(r2d_field), intent(inout), target :: fld
    call dlesm_check(dlesm_dev_field_alloc(int(size(fld%data,1), c_int), &
                                           int(size(fld%data,2), c_int), fld%device_ptr), &
                     'field_to_device: alloc')
    ! pin fld%data so the non-blocking hook copies are asynchronous
    call dlesm_check(dlesm_host_register(c_loc(fld%data), &
                     int(size(fld%data), c_size_t) * 8_c_size_t), 'field_to_device: pin')
    fld%read_from_device_c => dlesm_read_from_device_c
    fld%write_to_device_c  => dlesm_write_to_device_c
    fld%data_on_device = .true.
    call fld%write_to_device()          ! whole array, blocking (defaults of :467-525)
  end subroutine field_to_device

  !> free_field (src/field_mod.f90:395-403) does not touch device_ptr: release it here.
  subroutine field_free_device(fld)
    type(r2d_field), intent(inout), target :: fld
    if (.not. fld%data_on_device) return
    call dlesm_check(dlesm_host_unregister(c_loc(fld%data)), 'field_free_device: unpin')
    call dlesm_check(dlesm_dev_field_free(fld%device_ptr), 'field_free_device')
    fld%device_ptr = c_null_ptr
    fld%data_on_device = .false.
    nullify(fld%read_from_device_c, fld%write_to_device_c)
  end subroutine field_free_device

  !> copy_field(field_in, field_out), whole array (src/field_mod.f90:1152-1174)
  subroutine dev_copy_field(field_in, field_out)
    type(r2d_field), intent(in)    :: field_in
    type(r2d_field), intent(inout) :: field_out
    call dlesm_check(dlesm_dev_copy_field(field_in%device_ptr, field_out%device_ptr), 'copy_field')
  end subroutine dev_copy_field

  !> copy_field(field, src, dest) (src/field_mod.f90:1179-1187)
  subroutine dev_copy_field_patch(field, src, dest)
    type(r2d_field), intent(inout) :: field
    type(region_type), intent(in)  :: src, dest
    call dlesm_check(dlesm_dev_copy_patch(field%device_ptr,                       &
         int(src%xstart,c_int), int(src%xstop,c_int), int(src%ystart,c_int), int(src%ystop,c_int), &
         int(dest%xstart,c_int), int(dest%xstop,c_int), int(dest%ystart,c_int), int(dest%ystop,c_int)), &
         'copy_field(patch)')
  end subroutine dev_copy_field_patch

  !> set_field (src/field_mod.f90:1191-1202)
  subroutine dev_set_field(fld, val)
    type(r2d_field), intent(inout) :: fld
    real(go_wp), intent(in) :: val
    call dlesm_check(dlesm_dev_set_field(fld%device_ptr, real(val, c_double)), 'set_field')
  end subroutine dev_set_field

  !> field_checksum (src/field_mod.f90:1209-1219,1289-1307) without the whole-field D2H
  !! that get_data() forces: reduce on the device, then global_sum over ranks.
  function dev_field_checksum(field) result(val)
    type(r2d_field), intent(in) :: field
    real(go_wp) :: val
    real(c_double) :: s
    call dlesm_check(dlesm_dev_checksum(field%device_ptr,                              &
         int(field%internal%xstart,c_int), int(field%internal%xstop,c_int),             &
         int(field%internal%ystart,c_int), int(field%internal%ystop,c_int), s), 'field_checksum')
    call dlesm_check(dlesm_global_sum(s), 'field_checksum: global_sum')
    val = s
  end function dev_field_checksum

  !> fld%halo_exchange(depth) (src/field_mod.f90:1231-1256) entirely on the device: no
  !! read_halo_from_device / write_halo_to_device staging.  Tables: install_comms (below).
  subroutine dev_halo_exchange(fld)
    type(r2d_field), intent(inout) :: fld
    type(c_ptr) :: ptrs(1)
    ptrs(1) = fld%device_ptr
    call dlesm_check(dlesm_dev_halo_exchange(ptrs, 1_c_int), 'halo_exchange')
  end subroutine dev_halo_exchange

  !> Three fields of one phase in one grouped exchange (one message per neighbour)
  subroutine dev_halo_exchange_3(f1, f2, f3)
    type(r2d_field), intent(inout) :: f1, f2, f3
    type(c_ptr) :: ptrs(3)
    ptrs(1) = f1%device_ptr; ptrs(2) = f2%device_ptr; ptrs(3) = f3%device_ptr
    call dlesm_check(dlesm_dev_halo_exchange(ptrs, 3_c_int), 'halo_exchange(3)')
  end subroutine dev_halo_exchange_3

  !> Serial periodic BCs: the application's loop over fld%halo(:) calling copy_field
  !! (src/field_mod.f90:1394-1464) as one launch.
  subroutine dev_periodic_halos(fld)
    type(r2d_field), intent(inout) :: fld
    integer(c_int) :: px, py
    px = 0; py = 0
    if (fld%grid%boundary_conditions(1) == GO_BC_PERIODIC) px = 1
    if (fld%grid%boundary_conditions(2) == GO_BC_PERIODIC) py = 1
    call dlesm_check(dlesm_dev_periodic_halos(fld%device_ptr,                           &
         int(fld%internal%xstart,c_int), int(fld%internal%xstop,c_int),                  &
         int(fld%internal%ystart,c_int), int(fld%internal%ystop,c_int), px, py), 'periodic halos')
  end subroutine dev_periodic_halos

end module field_device_mod
