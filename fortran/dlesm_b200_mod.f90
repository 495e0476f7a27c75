!> ISO_C_BINDING interfaces to the device layer of libdlesm_b200.so (include/dlesm_b200.h).
!!
!! This is the whole Fortran <-> CUDA seam: scalars by value, arrays as type(c_ptr), one
!! interoperable derived type (dlesm_comms) that carries the map_comms tables of
!! src/parallel_comms_mod.f90:70-83.  No array descriptors and none of the reference's
!! derived types cross the boundary.
!!
!! SOURCE ONLY: neither the build container nor the B200 box has a Fortran compiler
!! (profiles/r01a_probe.txt), so this file has never been compiled here.  The identical call
!! sequence is exercised from C++/Python by tests/ (see INTEGRATION.md).
module dlesm_b200_mod
  use iso_c_binding
  implicit none
  public

  integer(c_int), parameter :: DLESM_OK = 0
  integer, parameter :: DLESM_MAXCOMM = 16   ! == MaxComm, src/parallel_comms_mod.f90:70

  !> Mirror of `dlesm_comms` (include/dlesm_b200.h); member order is the C struct's.
  type, bind(C) :: dlesm_comms
     integer(c_int) :: nsend, nrecv
     integer(c_int), dimension(DLESM_MAXCOMM) :: dirsend, destination
     integer(c_int), dimension(DLESM_MAXCOMM) :: isrcsend, jsrcsend
     integer(c_int), dimension(DLESM_MAXCOMM) :: idessend, jdessend
     integer(c_int), dimension(DLESM_MAXCOMM) :: nxsend, nysend, nsendp
     integer(c_int), dimension(DLESM_MAXCOMM) :: dirrecv, source
     integer(c_int), dimension(DLESM_MAXCOMM) :: isrcrecv, jsrcrecv
     integer(c_int), dimension(DLESM_MAXCOMM) :: idesrecv, jdesrecv
     integer(c_int), dimension(DLESM_MAXCOMM) :: nxrecv, nyrecv, nrecvp
  end type dlesm_comms

  interface
     ! ---- lifecycle (replaces parallel_utils_mod, src/parallel/parallel_utils_mod.f90:77-125)
     function dlesm_parallel_init(rank1, nranks) bind(C, name="dlesm_parallel_init") result(ierr)
       import :: c_int
       integer(c_int), value :: rank1, nranks
       integer(c_int) :: ierr
     end function
     function dlesm_device_init(device) bind(C, name="dlesm_device_init") result(ierr)
       import :: c_int
       integer(c_int), value :: device
       integer(c_int) :: ierr
     end function
     function dlesm_comm_unique_id(id128) bind(C, name="dlesm_comm_unique_id") result(ierr)
       import :: c_int, c_char
       character(kind=c_char), dimension(128), intent(out) :: id128
       integer(c_int) :: ierr
     end function
     function dlesm_comm_init(id128) bind(C, name="dlesm_comm_init") result(ierr)
       import :: c_int, c_char
       character(kind=c_char), dimension(128), intent(in) :: id128
       integer(c_int) :: ierr
     end function
     function dlesm_parallel_finalise() bind(C, name="dlesm_parallel_finalise") result(ierr)
       import :: c_int
       integer(c_int) :: ierr
     end function
     subroutine dlesm_parallel_abort(msg) bind(C, name="dlesm_parallel_abort")
       import :: c_char
       character(kind=c_char), dimension(*), intent(in) :: msg
     end subroutine
     function dlesm_global_sum(var) bind(C, name="dlesm_global_sum") result(ierr)
       import :: c_int, c_double
       real(c_double), intent(inout) :: var
       integer(c_int) :: ierr
     end function
     function dlesm_last_error() bind(C, name="dlesm_last_error") result(msg)
       import :: c_ptr
       type(c_ptr) :: msg
     end function

     ! ---- storage and the data_on_device hooks (src/field_mod.f90:65-105,128-134,153)
     function dlesm_dev_field_alloc(nx, ny, device_ptr) bind(C, name="dlesm_dev_field_alloc") result(ierr)
       import :: c_int, c_ptr
       integer(c_int), value :: nx, ny
       type(c_ptr), intent(out) :: device_ptr
       integer(c_int) :: ierr
     end function
     function dlesm_dev_field_free(device_ptr) bind(C, name="dlesm_dev_field_free") result(ierr)
       import :: c_int, c_ptr
       type(c_ptr), value :: device_ptr
       integer(c_int) :: ierr
     end function
     !> Exactly read_from_device_c_interface (src/field_mod.f90:66-72)
     subroutine dlesm_read_from_device_c(from, to, startx, starty, nx, ny, blocking) &
          bind(C, name="dlesm_read_from_device_c")
       import :: c_ptr, c_int, c_bool
       type(c_ptr), intent(in), value :: from
       type(c_ptr), intent(in), value :: to
       integer(c_int), intent(in), value :: startx, starty, nx, ny
       logical(c_bool), intent(in), value :: blocking
     end subroutine
     !> Exactly write_to_device_c_interface (src/field_mod.f90:87-93)
     subroutine dlesm_write_to_device_c(from, to, startx, starty, nx, ny, blocking) &
          bind(C, name="dlesm_write_to_device_c")
       import :: c_ptr, c_int, c_bool
       type(c_ptr), intent(in), value :: from
       type(c_ptr), intent(in), value :: to
       integer(c_int), intent(in), value :: startx, starty, nx, ny
       logical(c_bool), intent(in), value :: blocking
     end subroutine
     function dlesm_dev_sync() bind(C, name="dlesm_dev_sync") result(ierr)
       import :: c_int
       integer(c_int) :: ierr
     end function
     !> Pin an existing allocatable (fld%data) so that hook copies are truly asynchronous
     function dlesm_host_register(host_ptr, bytes) bind(C, name="dlesm_host_register") result(ierr)
       import :: c_int, c_ptr, c_size_t
       type(c_ptr), value :: host_ptr
       integer(c_size_t), value :: bytes
       integer(c_int) :: ierr
     end function
     function dlesm_host_unregister(host_ptr) bind(C, name="dlesm_host_unregister") result(ierr)
       import :: c_int, c_ptr
       type(c_ptr), value :: host_ptr
       integer(c_int) :: ierr
     end function
  end interface

  interface
     ! ---- field operations on device pointers (src/field_mod.f90:1126-1307)
     function dlesm_dev_copy_field(field_in, field_out) bind(C, name="dlesm_dev_copy_field") result(ierr)
       import :: c_int, c_ptr
       type(c_ptr), value :: field_in, field_out
       integer(c_int) :: ierr
     end function
     function dlesm_dev_copy_patch(field, sxstart, sxstop, systart, systop, &
                                   dxstart, dxstop, dystart, dystop)          &
          bind(C, name="dlesm_dev_copy_patch") result(ierr)
       import :: c_int, c_ptr
       type(c_ptr), value :: field
       integer(c_int), value :: sxstart, sxstop, systart, systop
       integer(c_int), value :: dxstart, dxstop, dystart, dystop
       integer(c_int) :: ierr
     end function
     function dlesm_dev_set_field(field, val) bind(C, name="dlesm_dev_set_field") result(ierr)
       import :: c_int, c_ptr, c_double
       type(c_ptr), value :: field
       real(c_double), value :: val
       integer(c_int) :: ierr
     end function
     function dlesm_dev_checksum(field, xstart, xstop, ystart, ystop, val) &
          bind(C, name="dlesm_dev_checksum") result(ierr)
       import :: c_int, c_ptr, c_double
       type(c_ptr), value :: field
       integer(c_int), value :: xstart, xstop, ystart, ystop
       real(c_double), intent(out) :: val
       integer(c_int) :: ierr
     end function
     function dlesm_dev_periodic_halos(field, xstart, xstop, ystart, ystop, periodic_x, periodic_y) &
          bind(C, name="dlesm_dev_periodic_halos") result(ierr)
       import :: c_int, c_ptr
       type(c_ptr), value :: field
       integer(c_int), value :: xstart, xstop, ystart, ystop, periodic_x, periodic_y
       integer(c_int) :: ierr
     end function

     ! ---- halo exchange (src/parallel_comms_mod.f90:1501-1855)
     function dlesm_dev_set_comms(tables) bind(C, name="dlesm_dev_set_comms") result(ierr)
       import :: c_int, dlesm_comms
       type(dlesm_comms), intent(in) :: tables
       integer(c_int) :: ierr
     end function
     function dlesm_dev_halo_exchange(fields, nfields) bind(C, name="dlesm_dev_halo_exchange") result(ierr)
       import :: c_int, c_ptr
       type(c_ptr), dimension(*), intent(in) :: fields
       integer(c_int), value :: nfields
       integer(c_int) :: ierr
     end function
     ! fld%halo_exchange of a host-resident field: patches up, device exchange, patches back
     function dlesm_host_halo_exchange(host, nx, ny) bind(C, name="dlesm_host_halo_exchange") result(ierr)
       import :: c_int, c_ptr
       type(c_ptr), value :: host
       integer(c_int), value :: nx, ny
       integer(c_int) :: ierr
     end function
     ! do_tile + GOCEAN_OMP_GRID="AxB" (src/field_mod.f90:300-325) -> block grid of the device sweeps
     function dlesm_dev_set_sweep_tiles(ntilex, ntiley) bind(C, name="dlesm_dev_set_sweep_tiles") result(ierr)
       import :: c_int
       integer(c_int), value :: ntilex, ntiley
       integer(c_int) :: ierr
     end function

     ! ---- grid upload (fills the 17 *_device slots of src/grid_mod.f90:104-150)
     function dlesm_dev_grid_upload(nx, ny, tmask, metrics, device_ptrs) &
          bind(C, name="dlesm_dev_grid_upload") result(ierr)
       import :: c_int, c_ptr
       integer(c_int), value :: nx, ny
       type(c_ptr), value :: tmask                         ! C_LOC(grid%tmask) or C_NULL_PTR
       type(c_ptr), dimension(16), intent(in) :: metrics   ! C_LOC of each metric array
       type(c_ptr), dimension(17), intent(out) :: device_ptrs
       integer(c_int) :: ierr
     end function
     !> frees the 17 values dlesm_dev_grid_upload returned (null entries skipped) and nulls them
     function dlesm_dev_grid_release(device_ptrs) bind(C, name="dlesm_dev_grid_release") result(ierr)
       import :: c_int, c_ptr
       type(c_ptr), dimension(17), intent(inout) :: device_ptrs
       integer(c_int) :: ierr
     end function

     ! ---- shallow-water stencils: one launcher per kernel + the fused step (SURVEY a-S)
     function dlesm_compute_cu(cu, p, u, xstart, xstop, ystart, ystop, tmask) &
          bind(C, name="dlesm_compute_cu") result(ierr)
       import :: c_int, c_ptr
       type(c_ptr), value :: cu, p, u, tmask
       integer(c_int), value :: xstart, xstop, ystart, ystop
       integer(c_int) :: ierr
     end function
     function dlesm_compute_cv(cv, p, v, xstart, xstop, ystart, ystop, tmask) &
          bind(C, name="dlesm_compute_cv") result(ierr)
       import :: c_int, c_ptr
       type(c_ptr), value :: cv, p, v, tmask
       integer(c_int), value :: xstart, xstop, ystart, ystop
       integer(c_int) :: ierr
     end function
     function dlesm_compute_z(z, p, u, v, dx, dy, xstart, xstop, ystart, ystop, tmask) &
          bind(C, name="dlesm_compute_z") result(ierr)
       import :: c_int, c_ptr, c_double
       type(c_ptr), value :: z, p, u, v, tmask
       real(c_double), value :: dx, dy
       integer(c_int), value :: xstart, xstop, ystart, ystop
       integer(c_int) :: ierr
     end function
     function dlesm_compute_h(h, p, u, v, xstart, xstop, ystart, ystop, tmask) &
          bind(C, name="dlesm_compute_h") result(ierr)
       import :: c_int, c_ptr
       type(c_ptr), value :: h, p, u, v, tmask
       integer(c_int), value :: xstart, xstop, ystart, ystop
       integer(c_int) :: ierr
     end function
     function dlesm_compute_unew(unew, uold, z, cv, h, tdt, dx, xstart, xstop, ystart, ystop, tmask, ucur) &
          bind(C, name="dlesm_compute_unew") result(ierr)
       import :: c_int, c_ptr, c_double
       type(c_ptr), value :: unew, uold, z, cv, h, tmask, ucur
       real(c_double), value :: tdt, dx
       integer(c_int), value :: xstart, xstop, ystart, ystop
       integer(c_int) :: ierr
     end function
     function dlesm_compute_vnew(vnew, vold, z, cu, h, tdt, dy, xstart, xstop, ystart, ystop, tmask, vcur) &
          bind(C, name="dlesm_compute_vnew") result(ierr)
       import :: c_int, c_ptr, c_double
       type(c_ptr), value :: vnew, vold, z, cu, h, tmask, vcur
       real(c_double), value :: tdt, dy
       integer(c_int), value :: xstart, xstop, ystart, ystop
       integer(c_int) :: ierr
     end function
     function dlesm_compute_pnew(pnew, pold, cu, cv, tdt, dx, dy, xstart, xstop, ystart, ystop, tmask, pcur) &
          bind(C, name="dlesm_compute_pnew") result(ierr)
       import :: c_int, c_ptr, c_double
       type(c_ptr), value :: pnew, pold, cu, cv, tmask, pcur
       real(c_double), value :: tdt, dx, dy
       integer(c_int), value :: xstart, xstop, ystart, ystop
       integer(c_int) :: ierr
     end function
     function dlesm_time_smooth(f, fnew, fold, alpha, xstart, xstop, ystart, ystop, tmask, points) &
          bind(C, name="dlesm_time_smooth") result(ierr)
       import :: c_int, c_ptr, c_double
       type(c_ptr), value :: f, fnew, fold, tmask
       real(c_double), value :: alpha
       integer(c_int), value :: xstart, xstop, ystart, ystop, points
       integer(c_int) :: ierr
     end function
     function dlesm_sw_step_fused(u, v, p, uold, vold, pold, unew, vnew, pnew, dx, dy, tdt, alpha, &
                                  first, xstart, xstop, ystart, ystop, tmask, variant)             &
          bind(C, name="dlesm_sw_step_fused") result(ierr)
       import :: c_int, c_ptr, c_double
       type(c_ptr), value :: u, v, p, uold, vold, pold, unew, vnew, pnew, tmask
       real(c_double), value :: dx, dy, tdt, alpha
       integer(c_int), value :: first, xstart, xstop, ystart, ystop, variant
       integer(c_int) :: ierr
     end function
     ! two model steps in one sweep (tile decomposed with halo_width >= 2; include/dlesm_b200.h)
     function dlesm_sw_step2_fused(u, v, p, uold, vold, pold, u2, v2, p2, uold2, vold2, pold2, dx, dy, tdt, &
                                   alpha, xstart, xstop, ystart, ystop, tmask, variant)                       &
          bind(C, name="dlesm_sw_step2_fused") result(ierr)
       import :: c_int, c_ptr, c_double
       type(c_ptr), value :: u, v, p, uold, vold, pold, u2, v2, p2, uold2, vold2, pold2, tmask
       real(c_double), value :: dx, dy, tdt, alpha
       integer(c_int), value :: xstart, xstop, ystart, ystop, variant
       integer(c_int) :: ierr
     end function
     function dlesm_dev_periodic_halos_deep(f, xstart, xstop, ystart, ystop, depth)                          &
          bind(C, name="dlesm_dev_periodic_halos_deep") result(ierr)
       import :: c_int, c_ptr
       type(c_ptr), value :: f
       integer(c_int), value :: xstart, xstop, ystart, ystop, depth
       integer(c_int) :: ierr
     end function
  end interface

  !> go_arg across the boundary (include/dlesm_b200.h `dlesm_go_arg`): src/argument_mod.f90:57-61
  !! plus the element family (0 field, 1 scalar, 2 grid property), since the integer values of
  !! GO_CU, GO_TIME_STEP and GO_I_SCALAR overlap.
  type, bind(C) :: dlesm_go_arg
     integer(c_int) :: arg_intent, kind, element
     integer(c_int) :: stencil(3)
  end type dlesm_go_arg

  interface
     ! ---- PSy layer: metadata-driven kernel dispatch (src/kernel_mod.f90, src/argument_mod.f90)
     function dlesm_kernel_create(name, cuda_source, meta_args, nargs, iterates_over, index_offset) &
          bind(C, name="dlesm_kernel_create") result(kern)
       import :: c_ptr, c_char, c_int, dlesm_go_arg
       character(kind=c_char), dimension(*), intent(in) :: name, cuda_source   ! NUL-terminated
       type(dlesm_go_arg), dimension(*), intent(in) :: meta_args
       integer(c_int), value :: nargs, iterates_over, index_offset
       type(c_ptr) :: kern
     end function
     !> arrays = fld%device_ptr / grid%*_device of the field and grid-array arguments in metadata
     !! order; reals / ints = scalar arguments and scalar grid properties in metadata order
     function dlesm_kernel_launch_dev(kern, arrays, narrays, reals, nreals, ints, nints, &
                                      xstart, xstop, ystart, ystop)                      &
          bind(C, name="dlesm_kernel_launch_dev") result(ierr)
       import :: c_ptr, c_int, c_double
       type(c_ptr), value :: kern
       type(c_ptr), dimension(*), intent(in) :: arrays
       real(c_double), dimension(*), intent(in) :: reals
       integer(c_int), dimension(*), intent(in) :: ints
       integer(c_int), value :: narrays, nreals, nints, xstart, xstop, ystart, ystop
       integer(c_int) :: ierr
     end function
     subroutine dlesm_kernel_free(kern) bind(C, name="dlesm_kernel_free")
       import :: c_ptr
       type(c_ptr), value :: kern
     end subroutine
  end interface

contains

  !> Turn a non-zero status into the reference's only error path: gocean_stop
  !! (src/gocean_mod.F90:50-57) is parallel_abort(msg), and in a B200 build parallel_abort is
  !! dlesm_parallel_abort (fortran/parallel_utils_nccl_mod.f90).  Called directly: this module
  !! must not `use gocean_mod` (gocean_mod -> parallel_mod -> parallel_utils_mod -> this module).
  subroutine dlesm_check(ierr, where)
    integer(c_int), intent(in) :: ierr
    character(len=*), intent(in) :: where
    character(kind=c_char), dimension(:), pointer :: cmsg
    character(len=320) :: msg
    integer :: i, n
    if (ierr == DLESM_OK) return
    msg = where//': '
    n = len_trim(msg) + 1
    call c_f_pointer(dlesm_last_error(), cmsg, (/ 256 /))
    do i = 1, 256
       if (cmsg(i) == c_null_char .or. n + i > len(msg) - 1) exit
       msg(n+i:n+i) = cmsg(i)
    end do
    call dlesm_parallel_abort(trim(msg)//c_null_char)
  end subroutine dlesm_check

  !> What r2d_field%read_from_device_c points at.  The pointer component's abstract interface
  !! (read_from_device_c_interface, src/field_mod.f90:66-72) has all-value interoperable dummies
  !! but is NOT bind(C), and BIND is a characteristic of a procedure: a bind(C) procedure is not a
  !! valid target for that pointer.  These two wrappers have exactly the reference's interface (no
  !! bind(C)) and forward to the C entry points.
  subroutine dlesm_read_from_device_hook(from, to, startx, starty, nx, ny, blocking)
    type(c_ptr), intent(in), value :: from
    type(c_ptr), intent(in), value :: to
    integer(c_int), intent(in), value :: startx, starty, nx, ny
    logical(c_bool), intent(in), value :: blocking
    call dlesm_read_from_device_c(from, to, startx, starty, nx, ny, blocking)
  end subroutine dlesm_read_from_device_hook

  !> What r2d_field%write_to_device_c points at (write_to_device_c_interface,
  !! src/field_mod.f90:87-93); see dlesm_read_from_device_hook.
  subroutine dlesm_write_to_device_hook(from, to, startx, starty, nx, ny, blocking)
    type(c_ptr), intent(in), value :: from
    type(c_ptr), intent(in), value :: to
    integer(c_int), intent(in), value :: startx, starty, nx, ny
    logical(c_bool), intent(in), value :: blocking
    call dlesm_write_to_device_c(from, to, startx, starty, nx, ny, blocking)
  end subroutine dlesm_write_to_device_hook

end module dlesm_b200_mod
