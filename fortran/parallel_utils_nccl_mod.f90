!> Third build-time variant of `parallel_utils_mod` (the reference selects between
!! parallel/parallel_utils_mod.f90 [MPI] and parallel/parallel_utils_stub_mod.f90 [serial]
!! at link time, src/Makefile:47-55): one process per B200, NCCL over NVLink instead of MPI.
!! Same module name and public names, so parallel_mod / parallel_comms_mod / field_mod
!! compile against it unchanged.
!!
!! Bring-up without MPI: the launcher starts one process per GPU and exports
!!   DLESM_RANK (0-based), DLESM_NRANKS, DLESM_ID_FILE (a path all ranks can read).
!! Rank 1 creates the 128-byte NCCL id and writes it to the file; the others poll for it.
!!
!! The host-staged point-to-point routines of the MPI variant (post_send / post_receive /
!! msg_wait*) are not provided: with device-resident fields fld%halo_exchange is served by
!! dlesm_dev_halo_exchange (field_device_mod), which moves packed edges GPU to GPU.  Calling
!! them stops, exactly as the serial stub does (parallel_utils_stub_mod.f90:81-117).
!! SOURCE ONLY -- never compiled here (no Fortran compiler on either box).
module parallel_utils_mod
  use iso_c_binding
  use kind_params_mod, only: go_wp
  use dlesm_b200_mod
  implicit none
  private

  integer :: nranks = 1      !< number of processes (GPUs)
  integer :: rank = 1        !< 1-based, like parallel_utils_mod.f90:84
  integer, parameter :: MSG_UNDEFINED = -99
  integer, parameter :: MSG_REQUEST_NULL = 0
  logical, parameter :: DIST_MEM_ENABLED = .true.

  public parallel_init, parallel_finalise, parallel_abort
  public get_rank, get_num_ranks, get_max_tag, gather
  public msg_wait, msg_wait_all, post_receive, post_send, global_sum
  public MSG_UNDEFINED, MSG_REQUEST_NULL, DIST_MEM_ENABLED

contains

  subroutine parallel_init()
    character(len=256) :: val, idfile
    character(kind=c_char), dimension(128) :: id
    integer :: ierr, lun, rank0, ios, tries
    rank0 = 0; nranks = 1
    call get_environment_variable('DLESM_RANK', val, status=ierr)
    if (ierr == 0) read(val, *) rank0
    call get_environment_variable('DLESM_NRANKS', val, status=ierr)
    if (ierr == 0) read(val, *) nranks
    rank = rank0 + 1
    call dlesm_check(dlesm_parallel_init(int(rank, c_int), int(nranks, c_int)), 'parallel_init')
    call dlesm_check(dlesm_device_init(-1_c_int), 'parallel_init: device')   ! LOCAL_RANK
    if (nranks == 1) return
    call get_environment_variable('DLESM_ID_FILE', idfile, status=ierr)
    if (ierr /= 0) call parallel_abort('parallel_init: DLESM_ID_FILE is not set')
    if (rank == 1) then
       call dlesm_check(dlesm_comm_unique_id(id), 'parallel_init: unique id')
       open(newunit=lun, file=trim(idfile)//'.tmp', access='stream', form='unformatted', &
            status='replace')
       write(lun) id
       close(lun)
       call rename(trim(idfile)//'.tmp', trim(idfile))      ! atomic publish (GNU extension)
    else
       do tries = 1, 600
          open(newunit=lun, file=trim(idfile), access='stream', form='unformatted', &
               status='old', iostat=ios)
          if (ios == 0) exit
          call sleep(1)
       end do
       if (ios /= 0) call parallel_abort('parallel_init: no NCCL id file')
       read(lun) id
       close(lun)
    end if
    call dlesm_check(dlesm_comm_init(id), 'parallel_init: comm')
  end subroutine parallel_init

  subroutine parallel_finalise()
    call dlesm_check(dlesm_parallel_finalise(), 'parallel_finalise')
  end subroutine parallel_finalise

  subroutine parallel_abort(msg)
    character(len=*), intent(in) :: msg
    call dlesm_parallel_abort(trim(msg)//c_null_char)     ! stderr, ncclCommAbort, exit(1)
  end subroutine parallel_abort

  function get_rank()
    integer :: get_rank
    get_rank = rank
  end function get_rank

  function get_num_ranks() result(num)
    integer :: num
    num = nranks
  end function get_num_ranks

  integer function get_max_tag()
    get_max_tag = huge(1) / 2     ! NCCL has no tags; ordering is by stream
  end function get_max_tag

  subroutine global_sum(var)
    real(go_wp), intent(inout) :: var
    real(c_double) :: v
    v = var
    call dlesm_check(dlesm_global_sum(v), 'global_sum')
    var = v
  end subroutine global_sum

  !> gather (parallel_utils_mod.f90:242-255): every rank contributes send_buffer, rank 1
  !! receives them back to back.  Served by the library (NCCL send/recv to rank 1).
  subroutine gather(send_buffer, recv_buffer)
    real(go_wp), dimension(:), intent(in), target :: send_buffer
    real(go_wp), dimension(:), intent(out), target :: recv_buffer
    interface
       function dlesm_gather_host(sendbuf, n, recvbuf) bind(C, name="dlesm_gather_host") result(ierr)
         import :: c_int, c_ptr, c_long_long
         type(c_ptr), value :: sendbuf, recvbuf
         integer(c_long_long), value :: n
         integer(c_int) :: ierr
       end function
    end interface
    call dlesm_check(dlesm_gather_host(c_loc(send_buffer), int(size(send_buffer), c_long_long), &
                                       c_loc(recv_buffer)), 'gather')
  end subroutine gather

  subroutine msg_wait(nmsg, flags, irecv)
    integer, intent(in) :: nmsg
    integer, dimension(:), intent(inout) :: flags
    integer, intent(out) :: irecv
    irecv = MSG_UNDEFINED
    call parallel_abort('msg_wait: host-staged messages are replaced by dlesm_dev_halo_exchange')
  end subroutine msg_wait

  subroutine msg_wait_all(nmsg, flags)
    integer, intent(in) :: nmsg
    integer, dimension(:), intent(inout) :: flags
    call parallel_abort('msg_wait_all: host-staged messages are replaced by dlesm_dev_halo_exchange')
  end subroutine msg_wait_all

  subroutine post_receive(nrecv, source, tag, exch_flag, rbuff, ibuff)
    integer, intent(in) :: nrecv, source
    integer :: tag, exch_flag
    real(go_wp), dimension(:), optional, intent(inout) :: rbuff
    integer, dimension(:), optional, intent(inout) :: ibuff
    call parallel_abort('post_receive: host-staged messages are replaced by dlesm_dev_halo_exchange')
  end subroutine post_receive

  subroutine post_send(sendBuff, nsend, destination, tag, exch_flag)
    integer, intent(in) :: nsend, destination
    real(kind=go_wp), dimension(nsend), intent(in) :: sendBuff
    integer :: tag, exch_flag
    call parallel_abort('post_send: host-staged messages are replaced by dlesm_dev_halo_exchange')
  end subroutine post_send

end module parallel_utils_mod
