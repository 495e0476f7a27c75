"""ctypes binding of the CPU oracle (oracle/liboracle.so).  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
may import this module; the product package dl_esm_inf_b200 never does.

Arrays are numpy, Fortran order, shape (nx, ny), so that a[i-1, j-1] is the reference's
data(i, j) and the linear offset is (j-1)*nx + (i-1).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liboracle.so")

MAX_DOMAINS = 4096
MAXCOMM = 16

# enums (src/grid_mod.f90:45-69, src/field_mod.f90:47-52, src/parallel_comms_mod.f90:101-110)
GO_ARAKAWA_C, GO_ARAKAWA_B = 0, 1
GO_OFFSET_SW, GO_OFFSET_SE, GO_OFFSET_NW, GO_OFFSET_NE, GO_OFFSET_ANY = 0, 1, 2, 3, 4
GO_BC_PERIODIC, GO_BC_EXTERNAL, GO_BC_NONE = 0, 1, 2
GO_U_POINTS, GO_V_POINTS, GO_T_POINTS, GO_F_POINTS, GO_ALL_POINTS = 0, 1, 2, 3, 4
NONE, Iplus, Iminus, Jplus, Jminus, IplusJplus, IminusJminus, IplusJminus, IminusJplus = range(9)

ERR_ARGS, ERR_NRANKS, ERR_HALO_WIDTH = -1, -2, -3
ERR_ALIGN_DIGITS, ERR_ALIGN_VALUE, ERR_PBC_DISTMEM = -4, -5, -6
ERR_GRID_NAME, ERR_GRID_OFFSET, ERR_FIELD_STOP, ERR_POINTS = -7, -8, -9, -10
ERR_MAXCOMM, ERR_HALO_DEPTH = -12, -13


def build(force: bool = False) -> str:
    """Compile oracle/liboracle.so with the committed Makefile (gcc only)."""
    srcs = [os.path.join(_HERE, f) for f in ("dlesm_oracle.c", "dlesm_oracle_sw.c", "dlesm_oracle.h")]
    stale = (not os.path.exists(_LIB_PATH)) or any(
        os.path.getmtime(s) > os.path.getmtime(_LIB_PATH) for s in srcs)
    if force or stale:
        subprocess.run(["make", "-C", _HERE, "-B", "liboracle.so"], check=True,
                       stdout=subprocess.DEVNULL)
    return _LIB_PATH


class Region(C.Structure):
    _fields_ = [("nx", C.c_int), ("ny", C.c_int), ("xstart", C.c_int), ("xstop", C.c_int),
                ("ystart", C.c_int), ("ystop", C.c_int)]

    def astuple(self):
        return (self.nx, self.ny, self.xstart, self.xstop, self.ystart, self.ystop)

    def box(self):
        return (self.xstart, self.xstop, self.ystart, self.ystop)

    def __repr__(self):
        return "Region(nx=%d, ny=%d, x=%d:%d, y=%d:%d)" % self.astuple()


class Subdomain(C.Structure):
    _fields_ = [("glob", Region), ("internal", Region)]  # `global` is a Python keyword


class Decomp(C.Structure):
    _fields_ = [("global_nx", C.c_int), ("global_ny", C.c_int), ("nx", C.c_int), ("ny", C.c_int),
                ("ndomains", C.c_int), ("max_width", C.c_int), ("max_height", C.c_int),
                ("domperrank", C.c_int), ("nranks", C.c_int),
                ("subdomains", Subdomain * MAX_DOMAINS),
                ("proc_subdomains", C.c_int * MAX_DOMAINS)]


class GridDims(C.Structure):
    _fields_ = [("nx", C.c_int), ("ny", C.c_int), ("padding", C.c_int),
                ("dx", C.c_double), ("dy", C.c_double)]


class Halo(C.Structure):
    _fields_ = [("source", Region), ("dest", Region)]


class FieldBounds(C.Structure):
    _fields_ = [("defined_on", C.c_int), ("internal", Region), ("whole", Region),
                ("num_halos", C.c_int), ("halo", Halo * 4)]


_ia = C.c_int * MAXCOMM


class Comms(C.Structure):
    _fields_ = [("nsend", C.c_int), ("nrecv", C.c_int)] + [
        (n, _ia) for n in ("dirsend", "destination", "isrcsend", "jsrcsend", "idessend",
                           "jdessend", "nxsend", "nysend", "nsendp", "dirrecv", "source",
                           "isrcrecv", "jsrcrecv", "idesrecv", "jdesrecv", "nxrecv", "nyrecv",
                           "nrecvp")]

    def sends(self):
        return [dict(dir=self.dirsend[k], dest=self.destination[k], isrc=self.isrcsend[k],
                     jsrc=self.jsrcsend[k], ides=self.idessend[k], jdes=self.jdessend[k],
                     nx=self.nxsend[k], ny=self.nysend[k]) for k in range(self.nsend)]

    def recvs(self):
        return [dict(dir=self.dirrecv[k], src=self.source[k], isrc=self.isrcrecv[k],
                     jsrc=self.jsrcrecv[k], ides=self.idesrecv[k], jdes=self.jdesrecv[k],
                     nx=self.nxrecv[k], ny=self.nyrecv[k]) for k in range(self.nrecv)]


class SwParams(C.Structure):
    _fields_ = [("nx", C.c_int), ("ny", C.c_int), ("dx", C.c_double), ("dy", C.c_double),
                ("dt", C.c_double), ("tdt", C.c_double), ("alpha", C.c_double)]


_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int)
_lib = None


def _d(a):
    if a is None:
        return None
    assert a.dtype == np.float64 and a.flags.f_contiguous, "need float64 Fortran-order array"
    return a.ctypes.data_as(_dp)


def _i(a):
    if a is None:
        return None
    assert a.dtype == np.int32 and a.flags.f_contiguous, "need int32 Fortran-order array"
    return a.ctypes.data_as(_ip)


def lib():
    global _lib
    if _lib is not None:
        return _lib
    build()
    L = C.CDLL(_LIB_PATH)
    R = C.POINTER(Region)
    FB = C.POINTER(FieldBounds)
    L.orc_go_decompose.argtypes = [C.c_int] * 7 + [C.POINTER(Decomp)]
    L.orc_parse_alignment.argtypes = [C.c_char_p, _ip]
    L.orc_grid_constructor_check.argtypes = [C.c_int, C.c_int]
    L.orc_grid_dims_compute.argtypes = [C.POINTER(Subdomain), C.c_int, C.POINTER(GridDims)]
    L.orc_grid_init.argtypes = [C.POINTER(Subdomain), _ip, C.c_int, C.c_int, C.c_int, C.c_double,
                                C.c_double, _ip, C.c_int, C.POINTER(GridDims), _ip,
                                C.POINTER(_dp)]
    L.orc_set_field_bounds.argtypes = [C.POINTER(Subdomain), C.c_int, C.c_int, C.c_int, _ip,
                                       C.c_int, FB]
    L.orc_field_construct.argtypes = [C.POINTER(Subdomain), FB, C.c_int, C.c_int, _dp, C.c_int, _dp]
    L.orc_field_construct.restype = None
    L.orc_copy_field.argtypes = [_dp, _dp, C.c_int, C.c_int]
    L.orc_copy_field.restype = None
    L.orc_copy_patch.argtypes = [_dp, C.c_int, R, R]
    L.orc_copy_patch.restype = None
    L.orc_set_field.argtypes = [_dp, C.c_int, C.c_int, C.c_double]
    L.orc_set_field.restype = None
    for f in (L.orc_array_checksum, L.orc_array_checksum_ld):
        f.argtypes = [_dp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]
        f.restype = C.c_double
    L.orc_apply_periodic_halos.argtypes = [_dp, C.c_int, FB]
    L.orc_apply_periodic_halos.restype = None
    for f in (L.orc_write_to_device, L.orc_read_from_device):
        f.argtypes = [_dp, _dp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]
        f.restype = None
    L.orc_gather_inner_data.argtypes = [C.POINTER(Decomp), C.c_int, C.POINTER(_dp), _ip, FB, _dp]
    L.orc_iprocmap.argtypes = [C.POINTER(Decomp), C.c_int, C.c_int, C.c_int]
    L.orc_map_comms.argtypes = [C.POINTER(Decomp), C.c_int, C.c_int, C.c_int, _ip, C.POINTER(Comms)]
    L.orc_exchange_generic_all.argtypes = [C.c_int, C.POINTER(_dp), _ip, C.POINTER(Comms)] + [C.c_int] * 4
    L.orc_compute_cu.argtypes = [_dp, _dp, _dp, C.c_int, R, _ip]
    L.orc_compute_cv.argtypes = [_dp, _dp, _dp, C.c_int, R, _ip]
    L.orc_compute_z.argtypes = [_dp, _dp, _dp, _dp, C.c_int, R, C.c_double, C.c_double, _ip]
    L.orc_compute_h.argtypes = [_dp, _dp, _dp, _dp, C.c_int, R, _ip]
    L.orc_compute_unew.argtypes = [_dp] * 5 + [C.c_int, R, C.c_double, C.c_double, _ip, _dp]
    L.orc_compute_vnew.argtypes = [_dp] * 5 + [C.c_int, R, C.c_double, C.c_double, _ip, _dp]
    L.orc_compute_pnew.argtypes = [_dp] * 4 + [C.c_int, R, C.c_double, C.c_double, C.c_double, _ip, _dp]
    L.orc_time_smooth.argtypes = [_dp, _dp, _dp, C.c_int, R, C.c_double, _ip, C.c_int]
    for n in ("orc_compute_cu", "orc_compute_cv", "orc_compute_z", "orc_compute_h",
              "orc_compute_unew", "orc_compute_vnew", "orc_compute_pnew", "orc_time_smooth",
              "orc_sw_init", "orc_sw_step", "orc_reset_saved_state"):
        getattr(L, n).restype = None
    L.orc_sw_init.argtypes = [_dp, _dp, _dp, C.c_int, C.c_int, R, C.c_double, C.c_double, FB, FB, FB]
    L.orc_sw_step.argtypes = [C.POINTER(_dp), C.POINTER(SwParams), FB, FB, FB, FB, _ip, C.c_int]
    L.orc_sw_run_threads.argtypes = [C.c_int] * 5 + [_dp]
    L.orc_sw_run_threads.restype = C.c_double
    L.orc_sw_last_checksums_ld.argtypes = [_dp]
    L.orc_sw_last_checksums_ld.restype = None
    _lib = L
    return L


# ---------------------------------------------------------------------------------------
# Friendly wrappers
# ---------------------------------------------------------------------------------------

class OracleStop(RuntimeError):
    """Stands in for gocean_stop / parallel_abort (src/gocean_mod.F90:50-57)."""

    def __init__(self, code, where):
        super().__init__("%s stopped with code %d" % (where, code))
        self.code = code


def _chk(rc, where):
    if rc != 0:
        raise OracleStop(rc, where)


def reset_saved_state():
    lib().orc_reset_saved_state()


def go_decompose(domainx, domainy, ndomains=None, ndomainx=None, ndomainy=None,
                 halo_width=None, nranks=1) -> Decomp:
    d = Decomp()
    a = lambda v: -1 if v is None else int(v)  # noqa: E731
    _chk(lib().orc_go_decompose(domainx, domainy, a(ndomains), a(ndomainx), a(ndomainy),
                                a(halo_width), nranks, C.byref(d)), "go_decompose")
    return d


def parse_alignment(envval):
    out = C.c_int(0)
    b = None if envval is None else envval.encode()
    _chk(lib().orc_parse_alignment(b, C.byref(out)), "grid_init(DL_ESM_ALIGNMENT)")
    return out.value


METRIC_NAMES = ("dx_t", "dy_t", "dx_u", "dy_u", "dx_v", "dy_v", "dx_f", "dy_f", "area_t",
                "area_u", "area_v", "gphiu", "gphiv", "gphif", "xt", "yt")
TMASK_UNSET = -12345  # marker for cells grid_init leaves uninitialised


class Grid:
    """What grid_type holds after decompose + grid_init (src/grid_mod.f90:75-157)."""

    def __init__(self, name, bcs, offset):
        _chk(lib().orc_grid_constructor_check(name, offset), "grid_constructor")
        self.name, self.offset = name, offset
        self.bcs = (C.c_int * 3)(*bcs)
        self.decomp = None
        self.subdomain = None
        self.rank = self.nranks = None

    def decompose(self, domainx, domainy, ndomains=None, ndomainx=None, ndomainy=None,
                  halo_width=None, rank=1, nranks=1):
        self.decomp = go_decompose(domainx, domainy, ndomains, ndomainx, ndomainy, halo_width,
                                   nranks)
        self.rank, self.nranks = rank, nranks
        self.subdomain = self.decomp.subdomains[rank - 1]  # src/grid_mod.f90:207
        self.global_nx, self.global_ny = self.decomp.global_nx, self.decomp.global_ny

    def init(self, dx, dy, tmask=None, alignment=1, dist_mem_enabled=None, metrics=True):
        if dist_mem_enabled is None:
            dist_mem_enabled = self.nranks > 1
        dims = GridDims()
        _chk(lib().orc_grid_dims_compute(C.byref(self.subdomain), alignment, C.byref(dims)),
             "grid_init")
        self.nx, self.ny = dims.nx, dims.ny
        self.tmask = np.full((self.nx, self.ny), TMASK_UNSET, dtype=np.int32, order="F")
        ptrs = (_dp * 16)()
        self.metrics = {}
        if metrics:
            for k, n in enumerate(METRIC_NAMES):
                self.metrics[n] = np.full((self.nx, self.ny), np.nan, order="F")
                ptrs[k] = _d(self.metrics[n])
        tin, tin_nx = None, 0
        if tmask is not None:
            tin = np.asfortranarray(tmask, dtype=np.int32)
            tin_nx = tin.shape[0]
        _chk(lib().orc_grid_init(C.byref(self.subdomain), self.bcs, self.nranks,
                                 int(bool(dist_mem_enabled)), alignment, dx, dy, _i(tin), tin_nx,
                                 C.byref(dims), _i(self.tmask), ptrs), "grid_init")
        self.dx, self.dy = dims.dx, dims.dy
        self.padding = dims.padding
        self.comms = None
        if dist_mem_enabled:
            self.comms = map_comms(self.decomp, self.rank, self.nranks)
        return self


def field_bounds(grid: Grid, points) -> FieldBounds:
    fb = FieldBounds()
    _chk(lib().orc_set_field_bounds(C.byref(grid.subdomain), grid.nx, grid.ny, grid.offset,
                                    grid.bcs, points, C.byref(fb)), "r2d_field")
    return fb


class Field:
    """r2d_field: bounds + data(1:grid%nx, 1:grid%ny) (src/field_mod.f90:139-166, 242-390)."""

    def __init__(self, grid: Grid, points, init_global_data=None):
        self.grid = grid
        self.b = field_bounds(grid, points)
        self.internal, self.whole = self.b.internal, self.b.whole
        self.num_halos = self.b.num_halos
        self.data = np.empty((grid.nx, grid.ny), order="F")
        g = None if init_global_data is None else np.asfortranarray(init_global_data, dtype=np.float64)
        lib().orc_field_construct(C.byref(grid.subdomain), C.byref(self.b), grid.nx, grid.ny,
                                  _d(g), 0 if g is None else g.shape[0], _d(self.data))

    def checksum(self, long_double=False):
        f = lib().orc_array_checksum_ld if long_double else lib().orc_array_checksum
        i = self.internal
        return f(_d(self.data), self.grid.nx, i.xstart, i.xstop, i.ystart, i.ystop)

    def apply_periodic_halos(self):
        lib().orc_apply_periodic_halos(_d(self.data), self.grid.nx, C.byref(self.b))


def copy_field(a_in, a_out):
    lib().orc_copy_field(_d(a_in), _d(a_out), a_in.shape[0], a_in.shape[1])


def copy_patch(a, src: Region, dest: Region):
    lib().orc_copy_patch(_d(a), a.shape[0], C.byref(src), C.byref(dest))


def set_field(a, val):
    lib().orc_set_field(_d(a), a.shape[0], a.shape[1], float(val))


def array_checksum(a, box=None, long_double=False):
    f = lib().orc_array_checksum_ld if long_double else lib().orc_array_checksum
    if box is None:
        box = (1, a.shape[0], 1, a.shape[1])
    return f(_d(a), a.shape[0], *box)


def make_region(xstart, xstop, ystart, ystop) -> Region:
    return Region(xstop - xstart + 1, ystop - ystart + 1, xstart, xstop, ystart, ystop)


def write_to_device(host, device, startx=1, starty=1, nx=None, ny=None):
    nx = host.shape[0] if nx is None else nx
    ny = host.shape[1] if ny is None else ny
    lib().orc_write_to_device(_d(host), _d(device), host.shape[0], startx, starty, nx, ny)


def read_from_device(device, host, startx=1, starty=1, nx=None, ny=None):
    nx = host.shape[0] if nx is None else nx
    ny = host.shape[1] if ny is None else ny
    lib().orc_read_from_device(_d(device), _d(host), host.shape[0], startx, starty, nx, ny)


def map_comms(decomp: Decomp, irank, nranks, pbc=False, halo_depths=(1, 1)) -> Comms:
    c = Comms()
    hd = (C.c_int * 2)(*halo_depths)
    _chk(lib().orc_map_comms(C.byref(decomp), irank, nranks, int(pbc), hd, C.byref(c)), "map_comms")
    return c


def exchange_all(fields, comms, dirs=(Jplus, Jminus, Iplus, Iminus)):
    """halo_exchange of one field on every rank (src/field_mod.f90:1247-1248)."""
    n = len(fields)
    ptrs = (_dp * n)(*[_d(f) for f in fields])
    pitches = (C.c_int * n)(*[f.shape[0] for f in fields])
    carr = (Comms * n)(*comms)
    _chk(lib().orc_exchange_generic_all(n, ptrs, pitches, carr, *dirs), "exchange_generic")


def gather_inner_data(decomp, fields):
    """fields: list of oracle Field objects, one per rank; returns the master's global array."""
    n = len(fields)
    ptrs = (_dp * n)(*[_d(f.data) for f in fields])
    pitches = (C.c_int * n)(*[f.data.shape[0] for f in fields])
    barr = (FieldBounds * n)(*[f.b for f in fields])
    out = np.zeros((decomp.global_nx, decomp.global_ny), order="F")
    _chk(lib().orc_gather_inner_data(C.byref(decomp), n, ptrs, pitches, barr, _d(out)),
         "gather_inner_data")
    return out


class ShallowWater:
    """Serial periodic shallow-water model in the reference-style un-fused formulation."""

    NAMES = ("u", "v", "p", "uold", "vold", "pold", "unew", "vnew", "pnew", "cu", "cv", "z", "h")

    def __init__(self, n_x, n_y, dx=1.0e5, dy=1.0e5, dt=90.0, alpha=0.001, alignment=1,
                 tmask=None):
        reset_saved_state()
        g = Grid(GO_ARAKAWA_C, (GO_BC_PERIODIC, GO_BC_PERIODIC, GO_BC_NONE), GO_OFFSET_SW)
        g.decompose(n_x, n_y)
        g.init(dx, dy, tmask=tmask, alignment=alignment, dist_mem_enabled=False, metrics=False)
        self.grid = g
        self.fb = {k: field_bounds(g, pts) for k, pts in
                   (("u", GO_U_POINTS), ("v", GO_V_POINTS), ("t", GO_T_POINTS), ("f", GO_F_POINTS))}
        self.f = {n: np.zeros((g.nx, g.ny), order="F") for n in self.NAMES}
        self.dt, self.tdt, self.alpha = dt, dt, alpha
        self.ncycle = 0
        self.tmask = g.tmask if tmask is not None else None
        lib().orc_sw_init(_d(self.f["u"]), _d(self.f["v"]), _d(self.f["p"]), g.nx, g.ny,
                          C.byref(self.fb["t"].internal), dx, dy, C.byref(self.fb["u"]),
                          C.byref(self.fb["v"]), C.byref(self.fb["t"]))
        for a, b in (("u", "uold"), ("v", "vold"), ("p", "pold")):
            copy_field(self.f[a], self.f[b])

    def params(self):
        g = self.grid
        return SwParams(g.nx, g.ny, g.dx, g.dy, self.dt, self.tdt, self.alpha)

    def step(self):
        ptrs = (_dp * 13)(*[_d(self.f[n]) for n in self.NAMES])
        prm = self.params()
        first = self.ncycle == 0
        lib().orc_sw_step(ptrs, C.byref(prm), C.byref(self.fb["u"]), C.byref(self.fb["v"]),
                          C.byref(self.fb["t"]), C.byref(self.fb["f"]), _i(self.tmask), int(first))
        if first:
            self.tdt = self.tdt + self.tdt
        self.ncycle += 1

    def checksums(self):
        i = self.fb["t"].internal
        return tuple(array_checksum(self.f[n], i.box()) for n in ("p", "u", "v"))


def sw_run_threads(global_nx, global_ny, nthreads, nsteps, warmup=0):
    cks = np.zeros(3)
    secs = lib().orc_sw_run_threads(global_nx, global_ny, nthreads, nsteps, warmup, _d(cks))
    if secs < 0:
        raise RuntimeError("orc_sw_run_threads failed")
    return secs, tuple(cks)


def sw_last_checksums_ld():
    """(p, u, v) checksums of the last sw_run_threads call, accumulated in long double"""
    cks = np.zeros(3)
    lib().orc_sw_last_checksums_ld(_d(cks))
    return tuple(cks)
