!> What a B200 build of grid_mod adds at the end of grid_init (after the map_comms call,
!! src/grid_mod.f90:566): upload tmask and the metric arrays once, filling the 17 `*_device`
!! c_ptr slots the reference declares (src/grid_mod.f90:104-150) and nulls (:286-302) but never
!! fills; and hand this rank's map_comms tables (module variables of
!! src/parallel_comms_mod.f90:70-83, computed by the UNCHANGED Fortran map_comms) to the
!! device exchange.
!! SOURCE ONLY -- never compiled here (no Fortran compiler on either box).
module grid_device_mod
  use iso_c_binding
  use grid_mod, only: grid_type
  use dlesm_b200_mod
  implicit none
  private
  public :: grid_to_device, install_comms

contains

  subroutine grid_to_device(grid)
    type(grid_type), intent(inout), target :: grid
    type(c_ptr) :: metrics(16), dev(17), tmask_ptr
    ! order = DLESM_GRID_DX_T .. DLESM_GRID_YT of include/dlesm_b200.h
    metrics(1)  = c_loc(grid%dx_t);   metrics(2)  = c_loc(grid%dy_t)
    metrics(3)  = c_loc(grid%dx_u);   metrics(4)  = c_loc(grid%dy_u)
    metrics(5)  = c_loc(grid%dx_v);   metrics(6)  = c_loc(grid%dy_v)
    metrics(7)  = c_loc(grid%dx_f);   metrics(8)  = c_loc(grid%dy_f)
    metrics(9)  = c_loc(grid%area_t); metrics(10) = c_loc(grid%area_u)
    metrics(11) = c_loc(grid%area_v)
    metrics(12) = c_loc(grid%gphiu);  metrics(13) = c_loc(grid%gphiv)
    metrics(14) = c_loc(grid%gphif)
    metrics(15) = c_loc(grid%xt);     metrics(16) = c_loc(grid%yt)
    tmask_ptr = c_null_ptr
    if (allocated(grid%tmask)) tmask_ptr = c_loc(grid%tmask)
    call dlesm_check(dlesm_dev_grid_upload(int(grid%nx, c_int), int(grid%ny, c_int), tmask_ptr, &
                                           metrics, dev), 'grid_to_device')
    grid%tmask_device  = dev(1)
    grid%dx_t_device   = dev(2);  grid%dy_t_device   = dev(3)
    grid%dx_u_device   = dev(4);  grid%dy_u_device   = dev(5)
    grid%dx_v_device   = dev(6);  grid%dy_v_device   = dev(7)
    grid%dx_f_device   = dev(8);  grid%dy_f_device   = dev(9)
    grid%area_t_device = dev(10); grid%area_u_device = dev(11); grid%area_v_device = dev(12)
    grid%gphiu_device  = dev(13); grid%gphiv_device  = dev(14); grid%gphif_device  = dev(15)
    grid%xt_device     = dev(16); grid%yt_device     = dev(17)
    call install_comms()
  end subroutine grid_to_device

  !> Copy the module-level tables of parallel_comms_mod into the interoperable struct.
  !! Every table used here is already public (src/parallel_comms_mod.f90:156-160); isrcrecv and
  !! jsrcrecv are private there and are not needed by the device exchange (set to -999, the
  !! reference's "unset" value, :246-261).
  subroutine install_comms()
    use parallel_comms_mod, only: nsend, nrecv, dirsend, destination, isrcsend, jsrcsend, &
         idessend, jdessend, nxsend, nysend, nsendp, dirrecv, source,                      &
         idesrecv, jdesrecv, nxrecv, nyrecv, nrecvp
    use parallel_utils_mod, only: get_num_ranks
    type(dlesm_comms) :: t
    if (get_num_ranks() == 1) return          ! map_comms is a no-op when serial (:216)
    t%nsend = nsend;             t%nrecv = nrecv
    t%dirsend = dirsend;         t%destination = destination
    t%isrcsend = isrcsend;       t%jsrcsend = jsrcsend
    t%idessend = idessend;       t%jdessend = jdessend
    t%nxsend = nxsend;           t%nysend = nysend
    t%nsendp = nsendp(:, 1)
    t%dirrecv = dirrecv;         t%source = source
    t%isrcrecv = -999;           t%jsrcrecv = -999
    t%idesrecv = idesrecv;       t%jdesrecv = jdesrecv
    t%nxrecv = nxrecv;           t%nyrecv = nyrecv
    t%nrecvp = nrecvp(:, 1)
    call dlesm_check(dlesm_dev_set_comms(t), 'install_comms')
  end subroutine install_comms

end module grid_device_mod
