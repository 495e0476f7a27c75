!> Helpers an application (or a PSy layer) uses next to the B200 build of field_mod
!! (fortran/patches/field_mod.f90.patch).  The public generics of field_mod -- copy_field,
!! set_field, field_checksum, fld%halo_exchange, free_field, the r2d_field constructor --
!! already dispatch on fld%data_on_device there; what is left here is
!!   * moving a field's VALUES to the device and back (the constructor gives every field
!!     device storage and the hooks; data_on_device stays .false. until this is called, the
!!     convention of PSyclone's generated layers),
!!   * the grouped three-field exchange (one message per neighbour for three fields), and
!!   * the serial periodic halo fill as one launch.
!! SOURCE ONLY -- no Fortran compiler on the build or GPU box; never compiled here.  The same
!! call sequence is driven through the same C entry points by tests/c_driver.c (plain C, no
!! Python in between) on the GPU.
module field_device_mod
  use iso_c_binding
  use kind_params_mod, only: go_wp
  use field_mod, only: r2d_field
  use grid_mod, only: GO_BC_PERIODIC
  use dlesm_b200_mod
  implicit none
  private
  public :: field_to_device, field_to_host
  public :: dev_halo_exchange_3, dev_periodic_halos

contains

  !> Upload the host values and mark the field device-resident.  Afterwards fld%get_data(),
  !! read_from_device(), write_to_device() (src/field_mod.f90:407-559) and every public generic
  !! of field_mod work on the device copy.
  subroutine field_to_device(fld)
    type(r2d_field), intent(inout), target :: fld
    if (fld%data_on_device) return
    ! pin fld%data so that non-blocking hook copies are asynchronous
    call dlesm_check(dlesm_host_register(c_loc(fld%data), &
                     int(size(fld%data), c_size_t) * 8_c_size_t), 'field_to_device: pin')
    fld%data_on_device = .true.
    call fld%write_to_device()          ! whole array, blocking (defaults of :467-525)
  end subroutine field_to_device

  !> Bring the values back and mark the field host-resident again (device storage is kept;
  !! free_field releases it).
  subroutine field_to_host(fld)
    type(r2d_field), intent(inout), target :: fld
    if (.not. fld%data_on_device) return
    call fld%read_from_device()
    fld%data_on_device = .false.
    call dlesm_check(dlesm_host_unregister(c_loc(fld%data)), 'field_to_host: unpin')
  end subroutine field_to_host

  !> Three device-resident fields of one phase in one grouped exchange (one message per
  !! neighbour).  fld%halo_exchange(depth) is the single-field form.
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
