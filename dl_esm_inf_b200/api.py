"""Host-side mirror of the reference's public Fortran API (fd_lib / dm_fd_lib) on top of the
C ABI of libdlesm_b200.so.  Names, argument meaning, 1-based inclusive regions and error
behaviour follow the Fortran modules so that tests read like the reference's own test
programs.  All field data is device-resident (data_on_device is always true); `fld.data` is
a host mirror that is synchronised only through get_data / set_data / read_from_device /
write_to_device, exactly like the reference's hooks (src/field_mod.f90:407-559).

Nothing here computes on the CPU: every operation is a call into the CUDA library.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from ._lib import CommPlan, Comms, FieldInfo, GridInfo, Region, Subdomain

# --- constants, same values as the Fortran parameters --------------------------------------
GO_ARAKAWA_C, GO_ARAKAWA_B = 0, 1                                       # src/grid_mod.f90:45-46
GO_OFFSET_SW, GO_OFFSET_SE, GO_OFFSET_NW, GO_OFFSET_NE, GO_OFFSET_ANY = 0, 1, 2, 3, 4  # :52-60
GO_BC_PERIODIC, GO_BC_EXTERNAL, GO_BC_NONE = 0, 1, 2                     # :64-69
GO_U_POINTS, GO_V_POINTS, GO_T_POINTS, GO_F_POINTS, GO_ALL_POINTS = 0, 1, 2, 3, 4  # field_mod:47-52
NBOUNDARY = 1                                                            # field_mod.f90:227
HALO_WIDTH_X = HALO_WIDTH_Y = 1                                          # grid_mod.f90:72-73
NONE, Iplus, Iminus, Jplus, Jminus, IplusJplus, IminusJminus, IplusJminus, IminusJplus = range(9)
MaxComm = 16
EXT_PERIODIC_DISTMEM = 1
EXT_HALO_DEPTH = 2
EXT_DIRTY_HALOS = 4
ARITH_EXACT, ARITH_CONTRACTED = 0, 1

GRID_ARRAYS = ("tmask", "dx_t", "dy_t", "dx_u", "dy_u", "dx_v", "dy_v", "dx_f", "dy_f", "area_t",
               "area_u", "area_v", "gphiu", "gphiv", "gphif", "xt", "yt")
SW_FIELDS = ("u", "v", "p", "uold", "vold", "pold", "unew", "vnew", "pnew", "cu", "cv", "z", "h")
SW_UNFUSED, SW_FUSED = 0, 1
HALO_EXCHANGE, HALO_PUSH = 0, 1
FUSED_PUSH = 0x1000


class GoceanStop(RuntimeError):
    """gocean_stop / parallel_abort (src/gocean_mod.F90:50-57): the reference terminates the
    program; the mirror raises so that a test can observe the same decision."""

    def __init__(self, code, msg):
        super().__init__("gocean_stop: %s (status %d)" % (msg, code))
        self.code = code
        self.msg = msg


def _chk(rc):
    if rc != 0:
        raise GoceanStop(rc, _lib.lib().dlesm_last_error().decode())


def _chk_ptr(p):
    if not p:
        L = _lib.lib()
        raise GoceanStop(L.dlesm_last_status(), L.dlesm_last_error().decode())
    return p


def _opt(v):
    return -1 if v is None else int(v)


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


# --- gocean_mod / parallel_mod ----------------------------------------------------------------
def gocean_initialise(rank=1, nranks=1, device=None, comm_id=None):
    """gocean_initialise -> parallel_init (src/gocean_mod.F90:22-34, src/parallel_mod.f90:51-63).
    rank is 1-based.  device/comm_id are needed only for device work / more than one rank."""
    L = _lib.lib()
    _chk(L.dlesm_parallel_init(rank, nranks))
    L.dlesm_reset_saved_state()
    if device is not None:
        _chk(L.dlesm_device_init(device))
    if comm_id is not None:
        _chk(L.dlesm_comm_init(C.c_char_p(bytes(comm_id))))


def comm_unique_id() -> bytes:
    buf = C.create_string_buffer(128)
    _chk(_lib.lib().dlesm_comm_unique_id(buf))
    return buf.raw


def gocean_finalise():
    _chk(_lib.lib().dlesm_parallel_finalise())


def gocean_stop(msg):
    raise GoceanStop(-1, msg)


def get_rank():
    return _lib.lib().dlesm_get_rank()


def get_num_ranks():
    return _lib.lib().dlesm_get_num_ranks()


def on_master():
    return bool(_lib.lib().dlesm_on_master())


def set_extensions(flags):
    _chk(_lib.lib().dlesm_set_extensions(flags))


def global_sum(val):
    v = C.c_double(val)
    _chk(_lib.lib().dlesm_global_sum(C.byref(v)))
    return v.value


class decomposition_type:
    """src/decomposition_mod.f90:54-68"""

    def __init__(self, handle, owned):
        self._h, self._owned = handle, owned
        info = (C.c_int * 8)()
        _chk(_lib.lib().dlesm_decomp_info(handle, info))
        (self.global_nx, self.global_ny, self.nx, self.ny, self.ndomains, self.max_width,
         self.max_height, self.domperrank) = list(info)

    @property
    def subdomains(self):
        out = []
        for k in range(1, self.ndomains + 1):
            s = Subdomain()
            _chk(_lib.lib().dlesm_decomp_subdomain(self._h, k, C.byref(s)))
            out.append(s)
        return out

    def proc_subdomains(self, islot, irank):
        return _lib.lib().dlesm_decomp_proc_subdomain(self._h, islot, irank)

    def __del__(self):
        if getattr(self, "_owned", False) and self._h:
            _lib.lib().dlesm_decomp_free(self._h)
            self._h = None


def go_decompose(domainx, domainy, ndomains=None, ndomainx=None, ndomainy=None, halo_width=None):
    """src/parallel_mod.f90:70-332"""
    h = _lib.lib().dlesm_go_decompose(domainx, domainy, _opt(ndomains), _opt(ndomainx),
                                      _opt(ndomainy), _opt(halo_width))
    return decomposition_type(_chk_ptr(h), owned=True)


def map_comms(decomp: decomposition_type, irank=None, nranks=None, pbc=(False, False),
              halo_depths=(1, 1)) -> Comms:
    """src/parallel_comms_mod.f90:178-1172 for rank `irank` (default: this rank)."""
    c = Comms()
    _chk(_lib.lib().dlesm_map_comms(decomp._h, irank or get_rank(), nranks or get_num_ranks(),
                                    int(pbc[0]), int(pbc[1]), halo_depths[0], halo_depths[1],
                                    C.byref(c)))
    return c


def comm_plan(comms: Comms, irank=None) -> CommPlan:
    """the per-peer message plan the device exchange derives from a rank's map_comms tables"""
    p = CommPlan()
    _chk(_lib.lib().dlesm_comm_plan_build(C.byref(comms), irank or get_rank(), C.byref(p)))
    return p


def iprocmap(decomp: decomposition_type, ia, ja, nranks=None):
    return _lib.lib().dlesm_iprocmap(decomp._h, nranks or get_num_ranks(), ia, ja)


# --- grid_mod -------------------------------------------------------------------------------
class grid_type:
    """grid_type(grid_name, boundary_conditions, grid_offsets) -- src/grid_mod.f90:75-157, 217-304"""

    def __init__(self, grid_name=None, boundary_conditions=None, grid_offsets=None, _handle=None):
        if _handle is not None:              # a tile of a tile set: the set owns the handle
            self._h, self._borrowed = _handle, True
            return
        bcs = (C.c_int * 3)(*boundary_conditions)
        self._h = _chk_ptr(_lib.lib().dlesm_grid_create(grid_name, bcs, _opt(grid_offsets)))

    def decompose(self, domainx, domainy, ndomains=None, ndomainx=None, ndomainy=None,
                  halo_width=None):
        """src/grid_mod.f90:183-211"""
        _chk(_lib.lib().dlesm_grid_decompose(self._h, domainx, domainy, _opt(ndomains),
                                             _opt(ndomainx), _opt(ndomainy), _opt(halo_width)))

    @property
    def info(self) -> GridInfo:
        gi = GridInfo()
        _chk(_lib.lib().dlesm_grid_get_info(self._h, C.byref(gi)))
        return gi

    # the public components of the Fortran type
    name = property(lambda s: s.info.name)
    offset = property(lambda s: s.info.offset)
    global_nx = property(lambda s: s.info.global_nx)
    global_ny = property(lambda s: s.info.global_ny)
    nx = property(lambda s: s.info.nx)
    ny = property(lambda s: s.info.ny)
    dx = property(lambda s: s.info.dx)
    dy = property(lambda s: s.info.dy)
    boundary_conditions = property(lambda s: tuple(s.info.boundary_conditions))
    subdomain = property(lambda s: s.info.subdomain)

    @property
    def decomp(self):
        return decomposition_type(_lib.lib().dlesm_grid_decomp(self._h), owned=False)

    @property
    def comms(self) -> Comms:
        c = Comms()
        _chk(_lib.lib().dlesm_grid_comms(self._h, C.byref(c)))
        return c

    def host_array(self, name):
        which = GRID_ARRAYS.index(name)
        gi = self.info
        out = np.empty((gi.nx, gi.ny), dtype=np.int32 if which == 0 else np.float64, order="F")
        _chk(_lib.lib().dlesm_grid_host_array(self._h, which, out.ctypes.data_as(C.c_void_p)))
        return out

    def get_tmask(self):
        """src/grid_mod.f90:169-177"""
        return self.host_array("tmask")

    def device_array(self, name):
        """the *_device c_ptr of src/grid_mod.f90:104-150, materialised on first use"""
        p = C.c_void_p()
        _chk(_lib.lib().dlesm_grid_device_array(self._h, GRID_ARRAYS.index(name), C.byref(p)))
        return p.value

    def field_bounds(self, grid_points) -> FieldInfo:
        """set_field_bounds (src/field_mod.f90:563-624) without allocating a field"""
        fi = FieldInfo()
        _chk(_lib.lib().dlesm_grid_field_bounds(self._h, grid_points, C.byref(fi)))
        return fi

    def __getattr__(self, item):
        if item in GRID_ARRAYS:
            return self.host_array(item)
        if item.endswith("_device") and item[:-7] in GRID_ARRAYS:
            return self.device_array(item[:-7])
        raise AttributeError(item)

    def __del__(self):
        if getattr(self, "_h", None) and not getattr(self, "_borrowed", False):
            _lib.lib().dlesm_grid_free(self._h)
            self._h = None


def grid_init(grid: grid_type, dxarg, dyarg, tmask=None):
    """src/grid_mod.f90:330-570.  Reads DL_ESM_ALIGNMENT from the environment."""
    if tmask is None:
        _chk(_lib.lib().dlesm_grid_init(grid._h, dxarg, dyarg, None, 0, 0))
    else:
        t = np.asfortranarray(tmask, dtype=np.int32)
        _chk(_lib.lib().dlesm_grid_init(grid._h, dxarg, dyarg, t.ctypes.data_as(C.POINTER(C.c_int)),
                                        t.shape[0], t.shape[1]))


# --- field_mod ------------------------------------------------------------------------------
class r2d_field:
    """r2d_field(grid, grid_points [, do_tile] [, init_global_data]) -- src/field_mod.f90:139-166,
    242-390."""

    def __init__(self, grid: grid_type = None, grid_points=None, do_tile=False,
                 init_global_data=None, _handle=None, _owned=True):
        self.grid = grid
        self._owned = _owned
        if _handle is not None:
            self._h = _handle
        else:
            g = None
            if init_global_data is not None:
                g = np.asfortranarray(init_global_data, dtype=np.float64)
            self._h = _chk_ptr(_lib.lib().dlesm_r2d_field_create(
                grid._h, grid_points, int(bool(do_tile)), None if g is None else _dp(g),
                0 if g is None else g.shape[0], 0 if g is None else g.shape[1]))
        self._data = None

    @property
    def info(self) -> FieldInfo:
        fi = FieldInfo()
        _chk(_lib.lib().dlesm_field_get_info(self._h, C.byref(fi)))
        return fi

    defined_on = property(lambda s: s.info.defined_on)
    internal = property(lambda s: s.info.internal)
    whole = property(lambda s: s.info.whole)
    num_halos = property(lambda s: s.info.num_halos)
    ntiles = property(lambda s: s.info.ntiles)
    data_on_device = property(lambda s: bool(s.info.data_on_device))
    device_ptr = property(lambda s: s.info.device_ptr)

    @property
    def halo(self):
        fi = self.info
        return [fi.halo[k] for k in range(max(fi.num_halos, 0))]

    @property
    def data(self):
        """host mirror of fld%data(1:grid%nx, 1:grid%ny); NOT synchronised implicitly"""
        if self._data is None:
            fi = self.info
            self._data = np.zeros((fi.nx, fi.ny), order="F")
        return self._data

    def get_data(self):
        """src/field_mod.f90:530-542: bring the whole array to the host and return it"""
        self.read_from_device()
        return self.data

    def set_data(self, array):
        """src/field_mod.f90:546-559"""
        self.data[:, :] = array
        self.write_to_device()
        return 0

    def read_from_device(self, startx=1, starty=1, nx=None, ny=None, blocking=True):
        """src/field_mod.f90:407-465"""
        d = self.data
        nx = d.shape[0] if nx is None else nx
        ny = d.shape[1] if ny is None else ny
        _chk(_lib.lib().dlesm_field_read_from_device(self._h, _dp(d), startx, starty, nx, ny, blocking))

    def write_to_device(self, startx=1, starty=1, nx=None, ny=None, blocking=True):
        """src/field_mod.f90:467-525"""
        d = self.data
        nx = d.shape[0] if nx is None else nx
        ny = d.shape[1] if ny is None else ny
        _chk(_lib.lib().dlesm_field_write_to_device(self._h, _dp(d), startx, starty, nx, ny, blocking))

    def halo_exchange(self, depth=1):
        """src/field_mod.f90:1231-1256"""
        _chk(_lib.lib().dlesm_field_halo_exchange(self._h, depth))

    def tile(self, itile):
        """tile(itile)%internal, %whole (src/field_mod.f90:316-319), 1 <= itile <= ntiles"""
        a, b = Region(), Region()
        _chk(_lib.lib().dlesm_field_tile(self._h, itile, C.byref(a), C.byref(b)))
        return a, b

    def halo_is_dirty(self, depth=1):
        """halo_type%needs_update as a query (src/halo_mod.f90:10-40)"""
        return bool(_lib.lib().dlesm_field_halo_is_dirty(self._h, depth))

    def set_halo_dirty(self):
        _chk(_lib.lib().dlesm_field_set_halo_dirty(self._h))

    def apply_periodic_halos(self):
        """for ihalo = 1..num_halos: copy_field(fld, halo%source, halo%dest) (SURVEY CS-6)"""
        _chk(_lib.lib().dlesm_field_apply_periodic_halos(self._h))

    def gather_inner_data(self):
        """src/field_mod.f90:1313-1390; the result is meaningful on the master only"""
        g = self.grid
        out = np.zeros((g.global_nx, g.global_ny), order="F")
        _chk(_lib.lib().dlesm_field_gather_inner_data(self._h, _dp(out)))
        return out

    def free(self):
        if self._h and self._owned:
            _chk(_lib.lib().dlesm_free_field(self._h))
        self._h = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


def halo_exchange_fields(fields, depth=1):
    """several fields in one grouped exchange (one message per neighbour)"""
    arr = (C.c_void_p * len(fields))(*[f._h for f in fields])
    _chk(_lib.lib().dlesm_fields_halo_exchange(arr, len(fields), depth))


def copy_field(a, b, dest=None):
    """copy_field generic (src/field_mod.f90:171-174): copy_field(field_in, field_out) or
    copy_field(field, src_region, dest_region)."""
    if dest is None:
        _chk(_lib.lib().dlesm_copy_field(a._h, b._h))
    else:
        _chk(_lib.lib().dlesm_copy_field_patch(a._h, C.byref(b), C.byref(dest)))


def set_field(fld, val):
    """src/field_mod.f90:1191-1202"""
    _chk(_lib.lib().dlesm_set_field(fld._h, float(val)))


def field_checksum(field):
    """src/field_mod.f90:1209-1219 (+ global_sum :1305)"""
    v = C.c_double(0.0)
    _chk(_lib.lib().dlesm_field_checksum(field._h, C.byref(v)))
    return v.value


def free_field(fld):
    """src/field_mod.f90:395-403"""
    fld.free()


def make_region(xstart, xstop, ystart, ystop) -> Region:
    return Region(xstop - xstart + 1, ystop - ystart + 1, xstart, xstop, ystart, ystop)


EXCHANGE_DIRECT, EXCHANGE_NCCL = 0, 1


def set_exchange_mode(mode):
    """EXCHANGE_DIRECT (default): halo patches go straight into the neighbours' receive slots over NVLink
    peer memory; EXCHANGE_NCCL: pack -> ncclSend/ncclRecv -> unpack.  Same on every rank."""
    _chk(_lib.lib().dlesm_dev_set_exchange_mode(int(mode)))


def dev_sync():
    _chk(_lib.lib().dlesm_dev_sync())


def kernel_launch_count():
    return _lib.lib().dlesm_kernel_launch_count()


def set_arith(mode):
    """ARITH_EXACT (one rounding per operation, bit-identical to the oracle) or ARITH_CONTRACTED
    (FMA contraction in the fused sweeps; within 1e-12 relative of the exact results)"""
    _chk(_lib.lib().dlesm_set_arith(mode))


def get_arith():
    return _lib.lib().dlesm_get_arith()


def set_sweep_tiles(ntilex=0, ntiley=0):
    """the tile grid of the sweeps, what GOCEAN_OMP_GRID="AxB" sets (0 = the library's choice)"""
    _chk(_lib.lib().dlesm_dev_set_sweep_tiles(ntilex, ntiley))


# --- application layer: shallow water ------------------------------------------------------------
class shallow_water:
    """The GOcean "shallow" time-stepping driven over fd_lib fields (not part of the reference
    library: SURVEY.md F2).  mode SW_UNFUSED launches one kernel per reference-style sweep;
    SW_FUSED runs the whole step as one sweep with pointer rotation."""

    def __init__(self, grid: grid_type, dt=90.0, alpha=0.001, mode=SW_FUSED, use_tmask=False):
        self.grid = grid
        self.mode = mode
        self.ncycle_hint = 0      # callers' scratch (bench.py notes whether the start-up steps were taken)
        self._h = _chk_ptr(_lib.lib().dlesm_sw_create(grid._h, dt, alpha, mode, int(use_tmask)))

    def field(self, name) -> r2d_field:
        h = _lib.lib().dlesm_sw_field(self._h, SW_FIELDS.index(name))
        if not h:
            raise KeyError("field %s is not allocated in this mode" % name)
        return r2d_field(self.grid, _handle=h, _owned=False)

    def init_device(self):
        _chk(_lib.lib().dlesm_sw_init_device(self._h))

    def set_state(self, u, v, p):
        a = [np.asfortranarray(x, dtype=np.float64) for x in (u, v, p)]
        _chk(_lib.lib().dlesm_sw_set_state(self._h, _dp(a[0]), _dp(a[1]), _dp(a[2])))

    def set_variant(self, variant):
        _chk(_lib.lib().dlesm_sw_set_variant(self._h, variant))

    def set_halo_mode(self, mode):
        """HALO_EXCHANGE (periodic copies / pack-NCCL-unpack) or HALO_PUSH (the fused sweep stores
        the neighbours' halo cells itself over NVLink peer memory).  Collective."""
        _chk(_lib.lib().dlesm_sw_set_halo_mode(self._h, mode))

    def set_two_step(self, on=True, variant2=0):
        """two model steps per sweep (temporal blocking; needs grid.decompose(..., halo_width=2))"""
        _chk(_lib.lib().dlesm_sw_set_two_step(self._h, int(bool(on)), variant2))

    def step(self, nsteps=1):
        _chk(_lib.lib().dlesm_sw_step(self._h, nsteps))

    def step_host(self, host_in, host_out=None, band_rows=0, nsteps=1):
        """nsteps (1, 2, or on one rank any even number) model steps with HOST state: host_in = (u, v, p, uold, vold, pold)
        Fortran-ordered (nx, ny) float64 arrays (or raw addresses of such buffers); results land in
        host_out (default: in place).  Upload, sweep and download are pipelined in row bands;
        nsteps=2 sends the bands through the two-step sweep (set_two_step), halving the PCIe
        bytes per model step."""
        if host_out is None:
            host_out = host_in
        PD = C.POINTER(C.c_double)

        def ptr(a):
            if isinstance(a, np.ndarray):
                assert a.dtype == np.float64 and a.flags.f_contiguous
                return a.ctypes.data_as(PD)
            return C.cast(a, PD)

        ins = (PD * 6)(*[ptr(a) for a in host_in])
        outs = (PD * 6)(*[ptr(a) for a in host_out])
        _chk(_lib.lib().dlesm_sw_step_host_n(self._h, ins, outs, int(band_rows), int(nsteps)))

    def checksums(self):
        return tuple(field_checksum(self.field(n)) for n in ("p", "u", "v"))

    def launch_grid(self, two_step=False):
        """(grid.x, grid.y) of the one-step or two-step sweep of this model"""
        g = (C.c_int * 2)()
        _chk(_lib.lib().dlesm_sw_launch_grid(self._h, int(bool(two_step)), g))
        return g[0], g[1]

    def free(self):
        """Collective when the object pushes halos on several ranks: call it on every rank."""
        if self._h:
            _lib.lib().dlesm_sw_free(self._h)
            self._h = None

    def __del__(self):
        # a finaliser must never enter a collective: local release only
        try:
            if self._h:
                _lib.lib().dlesm_sw_free_local(self._h)
                self._h = None
        except Exception:
            pass


# --- several sub-domains per rank (extension; src/decomposition_mod.f90:64-67, --------------------
# --- src/parallel_mod.f90:141-152, the \TODO of src/parallel_comms_mod.f90:295) -----------------
class tileset:
    """go_decompose(ndomains > nranks): this rank owns proc_subdomains(:, rank).  One ordinary grid per
    owned sub-domain (`grids`; grid_init each, create fields / models on them as usual) and ONE halo
    exchange over all of them: same-rank neighbours by device copies, the others packed per peer rank."""

    def __init__(self, grid_name, boundary_conditions, grid_offsets, domainx, domainy, ndomains=None,
                 ndomainx=None, ndomainy=None, halo_width=None):
        bcs = (C.c_int * 3)(*boundary_conditions)
        L = _lib.lib()
        self._h = _chk_ptr(L.dlesm_tileset_create(grid_name, bcs, _opt(grid_offsets), domainx, domainy,
                                                  _opt(ndomains), _opt(ndomainx), _opt(ndomainy), _opt(halo_width)))
        self.ntiles = L.dlesm_tileset_ntiles(self._h)
        self.grids = [grid_type(_handle=L.dlesm_tileset_grid(self._h, k)) for k in range(self.ntiles)]
        self.tile_ids = [L.dlesm_tileset_tile_id(self._h, k) for k in range(self.ntiles)]
        self.decomp = decomposition_type(L.dlesm_tileset_decomp(self._h), False)

    def _handles(self, fields_by_tile):
        """[[field q of tile k for k in tiles] for q in fields] -> flat field-major handle array"""
        flat = [f._h for per_tile in fields_by_tile for f in per_tile]
        return (C.c_void_p * len(flat))(*flat)

    def halo_exchange(self, fields_by_tile, depth=1):
        """fields_by_tile[q][k] = field q on tile k; collective"""
        assert all(len(per_tile) == self.ntiles for per_tile in fields_by_tile)
        _chk(_lib.lib().dlesm_tileset_halo_exchange(self._h, self._handles(fields_by_tile), len(fields_by_tile), depth))

    def gather_inner_data(self, fields):
        """one field per tile -> the global array on rank 1 (None elsewhere); collective"""
        d = self.decomp
        out = np.zeros((d.global_nx, d.global_ny), order="F")
        _chk(_lib.lib().dlesm_tileset_gather_inner_data(self._h, self._handles([fields]), _dp(out)))
        return out if on_master() else None

    def free(self):
        if self._h:
            _lib.lib().dlesm_tileset_free(self._h)
            self._h = None
            for g in self.grids:
                g._h = None


def tileset_plan(decomp: decomposition_type, irank, nranks, pbc=(False, False), depth=1):
    """host only: rank `irank`'s exchange plan over its tiles, as an (n, 14) integer array
    (kind, peer0, src_tile, dst_tile, si0, sj0, di0, dj0, nx, ny, offset, dst_id, src_id, dir) + totals"""
    rows = np.zeros((512, 14), dtype=np.int64)
    n = C.c_int(0)
    tot = (C.c_longlong * 2)()
    _chk(_lib.lib().dlesm_tileset_plan(decomp._h, irank, nranks, int(pbc[0]), int(pbc[1]), depth,
                                       rows.ctypes.data_as(C.POINTER(C.c_longlong)), 512, C.byref(n), tot))
    return rows[:n.value].copy(), (tot[0], tot[1])


class shallow_water_tiles:
    """the shallow-water model over a tile set: one fused-mode model per owned tile, stepped together
    (one sweep per tile, then one exchange of unew, vnew, pnew over all tiles)"""

    def __init__(self, ts: tileset, dt=90.0, alpha=0.001, use_tmask=False):
        self.ts = ts
        self.models = [shallow_water(g, dt=dt, alpha=alpha, mode=SW_FUSED, use_tmask=use_tmask) for g in ts.grids]

    def _handles(self):
        return (C.c_void_p * len(self.models))(*[m._h for m in self.models])

    def init_device(self):
        for m in self.models:
            m.init_device()
        _chk(_lib.lib().dlesm_tileset_sw_sync_state(self.ts._h, self._handles()))

    def set_state(self, states):
        """states[k] = (u, v, p) local arrays of tile k (internal points; halos are exchanged here)"""
        for m, (u, v, p) in zip(self.models, states):
            m.set_state(u, v, p)
        _chk(_lib.lib().dlesm_tileset_sw_sync_state(self.ts._h, self._handles()))

    def step(self, nsteps=1):
        _chk(_lib.lib().dlesm_tileset_sw_step(self.ts._h, self._handles(), nsteps))

    def fields(self, name):
        return [m.field(name) for m in self.models]

    def gather(self, name):
        return self.ts.gather_inner_data(self.fields(name))

    def checksums(self):
        """global sums over every tile of every rank"""
        return tuple(sum(field_checksum(f) for f in self.fields(n)) for n in ("p", "u", "v"))

    def free(self):
        for m in self.models:
            m.free()
        self.models = []


# --- kernel_mod / argument_mod: metadata-driven kernel dispatch (the PSy layer) ---------------
GO_READ, GO_WRITE, GO_READWRITE, GO_INC, GO_MIN, GO_MAX, GO_SUM = range(7)   # argument_mod.f90:39-46
GO_R_SCALAR, GO_I_SCALAR = 0, 1                                               # :68
GO_EVERY = 1
GO_CU, GO_CV, GO_CT, GO_CF = 1, 2, 3, 4                                       # :71
(GO_TIME_STEP, GO_GRID_AREA_T, GO_GRID_AREA_U, GO_GRID_AREA_V, GO_GRID_MASK_T, GO_GRID_DX_T,
 GO_GRID_DX_U, GO_GRID_DX_V, GO_GRID_DY_T, GO_GRID_DY_U, GO_GRID_DY_V, GO_GRID_LAT_U, GO_GRID_LAT_V,
 GO_GRID_DX_CONST, GO_GRID_DY_CONST, GO_GRID_X_MIN_INDEX, GO_GRID_X_MAX_INDEX, GO_GRID_Y_MIN_INDEX,
 GO_GRID_Y_MAX_INDEX) = range(1, 20)                                          # :75-112
GO_INTERNAL_PTS, GO_EXTERNAL_PTS, GO_ALL_PTS = 0, 1, 2                        # kernel_mod.f90:35-37
GO_POINTWISE = (0, 10, 0)                                                     # go_stencil(000,010,000)
ARG_FIELD, ARG_SCALAR, ARG_GRID = 0, 1, 2


def go_arg(arg_intent, element, stencil=GO_POINTWISE, kind=None):
    """go_arg(intent, element[, stencil]) -- src/argument_mod.f90:57-61.  `kind` says which family
    the element belongs to (the Fortran names carry it; their integer values overlap); use
    field_arg / scalar_arg / grid_arg to have it filled in."""
    a = _lib.GoArg()
    a.arg_intent, a.element = arg_intent, element
    a.kind = ARG_FIELD if kind is None else kind
    a.stencil[:] = list(stencil)
    return a


def field_arg(arg_intent, point_type, stencil=GO_POINTWISE):
    return go_arg(arg_intent, point_type, stencil, ARG_FIELD)


def scalar_arg(element, arg_intent=GO_READ):
    return go_arg(arg_intent, element, GO_POINTWISE, ARG_SCALAR)


def grid_arg(prop):
    return go_arg(GO_READ, prop, GO_POINTWISE, ARG_GRID)


class kernel_type:
    """`type, extends(kernel_type)` of src/kernel_mod.f90:46-50 / src/infrastructure_mod.f90:13-26:
    meta_args, ITERATES_OVER, index_offset and the per-point code -- here a CUDA C++ device
    function `<name>_code(int ji, int jj, ...)` -- compiled into an sm_100a sweep by the library."""

    def __init__(self, name, code, meta_args, iterates_over=GO_INTERNAL_PTS, index_offset=GO_OFFSET_ANY):
        arr = (_lib.GoArg * len(meta_args))(*meta_args)
        self.name = name
        self._h = _chk_ptr(_lib.lib().dlesm_kernel_create(name.encode(), code.encode(), arr, len(meta_args),
                                                          iterates_over, index_offset))

    @property
    def generated_source(self):
        return _lib.lib().dlesm_kernel_generated_source(self._h).decode()

    def __call__(self, fields, rscalars=(), iscalars=()):
        """invoke(kernel(args...)): fields / real scalars / integer scalars in metadata order"""
        fa = (C.c_void_p * max(1, len(fields)))(*[f._h for f in fields])
        ra = (C.c_double * max(1, len(rscalars)))(*rscalars)
        ia = (C.c_int * max(1, len(iscalars)))(*iscalars)
        _chk(_lib.lib().dlesm_kernel_launch(self._h, fa, len(fields), ra, len(rscalars), ia, len(iscalars)))

    def launch_dev(self, arrays, reals, ints, bounds):
        aa = (C.c_void_p * max(1, len(arrays)))(*arrays)
        ra = (C.c_double * max(1, len(reals)))(*reals)
        ia = (C.c_int * max(1, len(ints)))(*ints)
        _chk(_lib.lib().dlesm_kernel_launch_dev(self._h, aa, len(arrays), ra, len(reals), ia, len(ints), *bounds))

    def free(self):
        if getattr(self, "_h", None):
            _lib.lib().dlesm_kernel_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass
