"""Multi-rank GPU parity worker: one process per GPU, launched by tests/test_multi_gpu.py (or by
hand) as

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P tests/dist_worker.py

Every rank drives its tile through the public API (device-resident fields, NCCL / NVLink halo
exchange) and compares with the CPU oracle, which simulates ALL ranks of the same decomposition
in-process.  The cases are the reference's own distributed tests:
example/model.f90, tests/dist_mem/test_halos.f90, test_gsum.f90, test_reduction.f90, plus the
periodic shallow-water extension against the serial oracle run.
Prints "DIST OK <n checks>" on rank 0 and exits 0, or raises.
"""
from __future__ import annotations

import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import dl_esm_inf_b200 as fd  # noqa: E402
from dl_esm_inf_b200 import _lib  # noqa: E402
from oracle import oracle as O  # noqa: E402

RANK = int(os.environ.get("RANK", 0))
WORLD = int(os.environ.get("WORLD_SIZE", 1))
LOCAL = int(os.environ.get("LOCAL_RANK", 0))
BCS_EXT = (fd.GO_BC_EXTERNAL, fd.GO_BC_EXTERNAL, fd.GO_BC_NONE)
BCS_PER = (fd.GO_BC_PERIODIC, fd.GO_BC_PERIODIC, fd.GO_BC_NONE)
POINTS = (fd.GO_U_POINTS, fd.GO_V_POINTS, fd.GO_T_POINTS, fd.GO_F_POINTS)
NCHECK = 0
COLLECT = False       # bench.py's parity leg: count mismatches instead of stopping at the first
FAILURES = []


def ok(cond, msg):
    global NCHECK
    NCHECK += 1
    if not cond:
        if COLLECT:
            FAILURES.append("rank %d: %s" % (RANK + 1, msg))
            return
        raise AssertionError("rank %d: %s" % (RANK + 1, msg))


def bring_up():
    torch.cuda.set_device(LOCAL)
    dist.init_process_group("nccl", device_id=torch.device("cuda", LOCAL))
    fd.gocean_initialise(RANK + 1, WORLD, device=LOCAL)
    idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
    if RANK == 0:
        idt = torch.frombuffer(bytearray(fd.comm_unique_id()), dtype=torch.uint8).cuda()
    dist.broadcast(idt, 0)
    fd.gocean_initialise(RANK + 1, WORLD, device=LOCAL, comm_id=bytes(idt.cpu().numpy().tobytes()))


def oracle_grids(nx, ny, offset, bcs, tmask=True, alignment=1):
    """the oracle's grid of every rank"""
    out = []
    for r in range(1, WORLD + 1):
        O.reset_saved_state()
        g = O.Grid(O.GO_ARAKAWA_C, bcs, offset)
        g.decompose(nx, ny, rank=r, nranks=WORLD)
        tm = None
        if tmask:
            tm = np.ones((g.subdomain.glob.nx, g.subdomain.glob.ny), dtype=np.int32, order="F")
        g.init(1.0, 1.0, tmask=tm, alignment=alignment)
        out.append(g)
    return out


def my_grid(nx, ny, offset, bcs, tmask=True, dx=1.0, dy=1.0):
    g = fd.grid_type(fd.GO_ARAKAWA_C, bcs, offset)
    g.decompose(nx, ny)
    tm = None
    if tmask:
        sub = g.subdomain
        tm = np.ones((sub.glob.nx, sub.glob.ny), dtype=np.int32, order="F")
    fd.grid_init(g, dx, dy, tm)
    return g


def hill(og, fb_internal, points, nx_alloc, ny_alloc):
    """tests/dist_mem/test_halos.f90:142-189: externals get a wrong-but-plausible value, the
    internal region 10000*xpos + ypos at the staggered global position"""
    a = np.full((nx_alloc, ny_alloc), -1000.0 - og.rank, order="F")
    sub = og.subdomain
    i = fb_internal
    for jj in range(i.ystart, i.ystop + 1):
        for ji in range(i.xstart, i.xstop + 1):
            xpos = float(sub.glob.xstart + ji - sub.internal.xstart)
            ypos = float(sub.glob.ystart + jj - sub.internal.ystart)
            if points == fd.GO_U_POINTS or points == fd.GO_F_POINTS:
                xpos += 0.5
            if points == fd.GO_V_POINTS or points == fd.GO_F_POINTS:
                ypos += 0.5
            a[ji - 1, jj - 1] = 10000.0 * xpos + ypos
    return a


def case_example_model():
    """example/model.f90:54-109 at np = WORLD: 4x10, NE offset, external BCs, fields = rank"""
    ogs = oracle_grids(4, 10, O.GO_OFFSET_NE, BCS_EXT)
    g = my_grid(4, 10, fd.GO_OFFSET_NE, BCS_EXT)
    want = 0.0
    ofs = [O.Field(og, O.GO_T_POINTS) for og in ogs]
    for og, of in zip(ogs, ofs):
        of.data[:, :] = float(og.rank)
        want += of.checksum()
    for pts in POINTS:
        f = fd.r2d_field(g, pts)
        fd.set_field(f, float(RANK + 1))
        f.halo_exchange(1)
        got = fd.field_checksum(f)
        ok(abs(got - want) <= 1e-12 * abs(want), "example checksum %r != %r" % (got, want))
        f.free()
    if WORLD == 2:
        ok(want == 60.0, "derived np=2 answer of example/model.f90 is 60.0, oracle says %r" % want)


def case_halos(jpiglo, jpjglo, alignment=None):
    """tests/dist_mem/test_halos.f90 pattern, all four staggerings, compared with the oracle's
    exchange_generic on every cell of the local array (edges and corners)"""
    if alignment:
        os.environ["DL_ESM_ALIGNMENT"] = str(alignment)
    else:
        os.environ.pop("DL_ESM_ALIGNMENT", None)
    ogs = oracle_grids(jpiglo, jpjglo, O.GO_OFFSET_NE, BCS_EXT, alignment=alignment or 1)
    g = my_grid(jpiglo, jpjglo, fd.GO_OFFSET_NE, BCS_EXT)
    og = ogs[RANK]
    ok((g.nx, g.ny) == (og.nx, og.ny), "allocation shape %r vs oracle %r" % ((g.nx, g.ny), (og.nx, og.ny)))
    fields, wants = [], []
    for pts in POINTS:
        arrs = []
        for o in ogs:
            fb = O.field_bounds(o, pts)
            arrs.append(hill(o, fb.internal, pts, o.nx, o.ny))
        mine = arrs[RANK].copy(order="F")
        if WORLD > 1:            # one rank: no tables, halo_exchange is a no-op (src/field_mod.f90:1231-1256)
            O.exchange_all(arrs, [o.comms for o in ogs])
        f = fd.r2d_field(g, pts)
        f.set_data(mine)
        fields.append(f)
        wants.append(arrs[RANK])
    # two fields one by one, two in one grouped exchange
    fields[0].halo_exchange(1)
    fields[1].halo_exchange(1)
    fd.halo_exchange_fields(fields[2:], 1)
    for pts, f, want in zip(POINTS, fields, wants):
        got = f.get_data()
        ok(np.array_equal(got, want), "halo exchange %dx%d points %d differs from the oracle at %s"
           % (jpiglo, jpjglo, pts, np.argwhere(got != want)[:4].tolist()))
        # the reference's own assertion: edge halos equal hill within 1e-8 (test_halos.f90:191-265)
        f.free()
    os.environ.pop("DL_ESM_ALIGNMENT", None)


def case_gsum():
    """tests/dist_mem/test_gsum.f90: internal = 1, external = -100, checksum = jpiglo*jpjglo"""
    g = my_grid(4, 10, fd.GO_OFFSET_NE, BCS_EXT)
    for pts in POINTS:
        f = fd.r2d_field(g, pts)
        a = np.full((g.nx, g.ny), -100.0, order="F")
        i = f.internal
        a[i.xstart - 1:i.xstop, i.ystart - 1:i.ystop] = 1.0
        f.set_data(a)
        got = fd.field_checksum(f)
        ok(abs(got - 40.0) < 1e-8, "gsum %r != 40" % got)
        f.free()


def case_reduction(n=10):
    """tests/dist_mem/test_reduction.f90: scatter (i-1)+(j-1)*n, +1, gather on the master"""
    g = fd.grid_type(fd.GO_ARAKAWA_C, BCS_EXT, fd.GO_OFFSET_NE)
    g.decompose(n, n)
    fd.grid_init(g, 1.0, 1.0)                      # no tmask, as the reference test
    glob = np.asfortranarray(np.arange(n)[:, None] + np.arange(n)[None, :] * float(n))
    f = fd.r2d_field(g, fd.GO_T_POINTS, init_global_data=glob)
    sub = g.subdomain
    a = f.get_data().copy(order="F")
    i = sub.internal
    for jj in range(i.ystart, i.ystop + 1):
        for ji in range(i.xstart, i.xstop + 1):
            gi, gj = sub.glob.xstart + ji - i.xstart, sub.glob.ystart + jj - i.ystart
            ok(a[ji - 1, jj - 1] == (gi - 1) + (gj - 1) * n, "scatter value at (%d,%d)" % (ji, jj))
    f.set_data(a + 1.0)
    out = f.gather_inner_data()
    if RANK == 0:
        ok(np.array_equal(out, glob + 1.0), "gather_inner_data differs")
    f.free()


def case_halos_depth2(n_x, n_y, depth=2):
    """Extension (BASELINE config 5 "width-2 halos"): grid decomposed with
    halo_width = depth, halo_exchange(depth).  Known answer, independent of any table: after the
    exchange every halo cell within `depth` of the internal region that lies inside the global
    domain holds the hill value of its global position (corners included)."""
    os.environ.pop("DL_ESM_ALIGNMENT", None)
    fd.set_extensions(fd.EXT_HALO_DEPTH)
    g = fd.grid_type(fd.GO_ARAKAWA_C, BCS_EXT, fd.GO_OFFSET_NE)
    g.decompose(n_x, n_y, halo_width=depth)
    sub = g.subdomain
    tm = np.ones((sub.glob.nx, sub.glob.ny), dtype=np.int32, order="F")
    fd.grid_init(g, 1.0, 1.0, tm)
    li = sub.internal
    ok(li.xstart == depth + 1 and li.ystart == depth + 1, "internal region starts after %d halo cells" % depth)
    flds, srcs = [], []
    for k, pts in enumerate(POINTS):
        f = fd.r2d_field(g, pts)
        a = np.full((g.nx, g.ny), -777.0 - RANK, order="F")
        ii = np.arange(li.xstart, li.xstop + 1)
        jj = np.arange(li.ystart, li.ystop + 1)
        a[np.ix_(ii - 1, jj - 1)] = (10000.0 * (sub.glob.xstart + ii - li.xstart)[:, None]
                                     + (sub.glob.ystart + jj - li.ystart)[None, :] + 0.25 * k)
        f.set_data(a)
        flds.append(f)
        srcs.append(a)
    flds[0].halo_exchange(depth)
    fd.halo_exchange_fields(flds[1:], depth)
    for k, (f, src) in enumerate(zip(flds, srcs)):
        got = f.get_data()
        want = src.copy(order="F")
        for ji in range(li.xstart - depth, li.xstop + depth + 1):
            for jj_ in range(li.ystart - depth, li.ystop + depth + 1):
                gx, gy = sub.glob.xstart + ji - li.xstart, sub.glob.ystart + jj_ - li.ystart
                if 1 <= gx <= n_x and 1 <= gy <= n_y:
                    want[ji - 1, jj_ - 1] = 10000.0 * gx + gy + 0.25 * k
        ok(np.array_equal(got, want), "depth-%d exchange %dx%d field %d differs at %s"
           % (depth, n_x, n_y, k, np.argwhere(got != want)[:4].tolist()))
        f.free()
    # without the extension the depth argument is ignored, as in the reference
    fd.set_extensions(0)
    f = fd.r2d_field(g, fd.GO_T_POINTS)
    f.set_data(srcs[2])
    f.halo_exchange(depth)
    got = f.get_data()
    outer = got.copy()
    outer[li.xstart - 2:li.xstop + 1, li.ystart - 2:li.ystop + 1] = srcs[2][li.xstart - 2:li.xstop + 1, li.ystart - 2:li.ystop + 1]
    ok(np.array_equal(outer, srcs[2]), "depth is ignored without the extension: only the width-1 ring may change")
    f.free()
    fd.gocean_initialise(RANK + 1, WORLD)           # resets go_decompose's SAVEd halo width to 1


def case_nemolite2d(n_x, n_y, nsteps, alignment):
    """The reference's own distributed configuration (NE offset, external BCs, T-mask) driving the
    NEMOLite2D-style application on the PSy layer: N tiles with halo_exchange between the kernels
    == the serial oracle of the whole basin, on the gathered internal fields."""
    from examples import nemolite2d as N
    from oracle import nemolite2d_oracle as NO
    if alignment:
        os.environ["DL_ESM_ALIGNMENT"] = str(alignment)
    else:
        os.environ.pop("DL_ESM_ALIGNMENT", None)
    dx, dy = 1000.0, 1200.0
    tm, state = N.global_setup(n_x, n_y)
    g = fd.grid_type(fd.GO_ARAKAWA_C, BCS_EXT, fd.GO_OFFSET_NE)
    g.decompose(n_x, n_y)
    fd.grid_init(g, dx, dy, N.local_tmask(g, tm))
    model = N.NemoLite2D(g)
    model.set_state(N.local_state(g, state))
    om = NO.Model.serial_global(n_x, n_y, tm, state, dx, dy, model.rdt, model.cbfr, model.visc, N.OMEGA, N.D2R, N.G)
    for step in range(1, nsteps + 1):
        model.step(step)
        om.step(step)
    for name in ("ssha", "sshn", "sshn_u", "sshn_v", "un", "vn", "ua", "va"):
        out = model.f[name].gather_inner_data()
        if RANK == 0:
            want = om.a[name][1:n_x + 1, 1:n_y + 1]
            err = np.abs(out - want)
            ok(bool(np.all(err <= 1e-12 * np.maximum(np.abs(want), 1e-3))),
               "nemolite2d %dx%d field %s: max abs err %.3e (sin() differs in the last bit between libm and the "
               "device; everything else is exact)" % (n_x, n_y, name, float(err.max())))
            ok(not np.array_equal(want, state[name][1:n_x + 1, 1:n_y + 1]) or name in ("sshn_u",),
               "field %s evolved" % name)
    model.free()
    os.environ.pop("DL_ESM_ALIGNMENT", None)


def case_gather():
    """parallel_utils_mod gather (src/parallel/parallel_utils_mod.f90:242-255) over NCCL"""
    import ctypes as C
    n = 37
    send = np.arange(n, dtype=np.float64) + 1000.0 * (RANK + 1)
    recv = np.zeros(n * WORLD)
    PD = C.POINTER(C.c_double)
    rc = _lib.lib().dlesm_gather_host(send.ctypes.data_as(PD), n, recv.ctypes.data_as(PD))
    ok(rc == 0, "gather rc %d" % rc)
    if RANK == 0:
        want = np.concatenate([np.arange(n) + 1000.0 * (r + 1) for r in range(WORLD)])
        ok(np.array_equal(recv, want), "gather differs")


def island_mask(n_x, n_y):
    """land islands on the global (n_x, n_y) T points (the pattern of tests/test_gpu_parity.py)"""
    i = np.arange(1, n_x + 1)[:, None]
    j = np.arange(1, n_y + 1)[None, :]
    land = ((i % 64 >= 20) & (i % 64 <= 27)) & ((j % 64 >= 30) & (j % 64 <= 37))
    return np.where(land, 0, 1).astype(np.int32)


def case_shallow_water(n_x, n_y, nsteps, alignment, mode, halo_mode=None, two_step=False, masked=False,
                       arith=None):
    """periodic shallow water across ranks (extension F4) == the serial oracle, bit for bit (with
    arith=ARITH_CONTRACTED: within 1e-12 of it on the scale of each field);
    halo_mode HALO_PUSH: the fused sweep stores the neighbours' halo cells itself (NVLink peer
    memory + flag hand-shake) instead of pack -> NCCL -> unpack"""
    os.environ["DL_ESM_ALIGNMENT"] = str(alignment)
    if arith is not None:
        fd.set_arith(arith)
    fd.set_extensions(fd.EXT_PERIODIC_DISTMEM | (fd.EXT_HALO_DEPTH if two_step else 0))
    gmask = island_mask(n_x, n_y) if masked else None
    osw = O.ShallowWater(n_x, n_y, alignment=1,
                         tmask=np.asfortranarray(np.pad(gmask, 1, mode="wrap")) if masked else None)
    init = {k: osw.f[k].copy(order="F") for k in ("u", "v", "p")}
    for _ in range(nsteps):
        osw.step()
    g = fd.grid_type(fd.GO_ARAKAWA_C, BCS_PER, fd.GO_OFFSET_SW)
    if two_step:
        # two model steps per sweep: tiles with two halo cells, halo_exchange(depth 2) of the six
        # prognostic fields every second step (both extensions)
        g.decompose(n_x, n_y, halo_width=2)
    else:
        g.decompose(n_x, n_y)
    tm = None
    if masked:
        # this tile's part of the global mask over its whole extent (halo cells wrap around)
        sb, hw = g.subdomain, g.subdomain.internal.xstart - 1
        mi = (sb.glob.xstart - 1 - hw + np.arange(sb.glob.nx)) % n_x
        mj = (sb.glob.ystart - 1 - hw + np.arange(sb.glob.ny)) % n_y
        tm = np.asfortranarray(gmask[np.ix_(mi, mj)])
    fd.grid_init(g, 1.0e5, 1.0e5, tm)
    sw = fd.shallow_water(g, mode=mode, use_tmask=masked)
    if two_step:
        sw.set_two_step(True)
        if halo_mode is not None:
            sw.set_halo_mode(halo_mode)              # two-step sweep pushing its own depth-2 halos
            halo_mode = None
    # this tile's part of the oracle's initial state, halo ring included (periodic wrap)
    sub = g.subdomain
    li = sub.internal
    ji = np.arange(li.xstart - 1, li.xstop + 2)           # local 1-based columns incl. halo
    jj = np.arange(li.ystart - 1, li.ystop + 2)
    gi = (sub.glob.xstart + (ji - li.xstart) - 1) % n_x + 1   # 0-based index into the oracle array
    gj = (sub.glob.ystart + (jj - li.ystart) - 1) % n_y + 1
    local = {}
    for k in ("u", "v", "p"):
        a = np.zeros((g.nx, g.ny), order="F")
        a[np.ix_(ji - 1, jj - 1)] = init[k][np.ix_(gi, gj)]
        local[k] = a
    sw.set_state(local["u"], local["v"], local["p"])
    if halo_mode is not None:
        sw.set_halo_mode(halo_mode)
    sw.step(nsteps)
    fd.dev_sync()                                   # also reports a push hand-shake time-out
    oi = osw.fb["t"].internal
    for name in ("u", "v", "p", "uold", "vold", "pold"):
        out = sw.field(name).gather_inner_data()
        if RANK == 0:
            want = osw.f[name][oi.xstart - 1:oi.xstop, oi.ystart - 1:oi.ystop]
            if arith == fd.ARITH_CONTRACTED:
                err = float(np.max(np.abs(out - want)) / np.max(np.abs(want)))
                ok(err <= 1e-12, "shallow water %dx%d contracted arithmetic, field %s: max |diff| / max |field| = %.3e"
                   % (n_x, n_y, name, err))
                continue
            ok(np.array_equal(out, want), "shallow water %dx%d mode %d field %s: %d cells differ, max rel %.3e"
               % (n_x, n_y, mode, name, int((out != want).sum()),
                  float(np.max(np.abs(out - want) / np.maximum(np.abs(want), 1e-300)))))
    got, want = sw.checksums(), osw.checksums()
    for a, b in zip(got, want):
        ok(abs(a - b) <= 1e-12 * abs(b), "checksum %r vs oracle %r" % (a, b))
    sw.free()
    fd.set_extensions(0)
    if arith is not None:
        fd.set_arith(fd.ARITH_EXACT)
    os.environ.pop("DL_ESM_ALIGNMENT", None)
    if two_step:
        fd.gocean_initialise(RANK + 1, WORLD)       # resets go_decompose's SAVEd halo width to 1


def case_tileset(n_x, n_y, tiles_per_rank, nsteps, alignment, masked=False):
    """extension F3: `tiles_per_rank` sub-domains per rank (go_decompose(ndomains > nranks),
    src/parallel_mod.f90:141-152; proc_subdomains, src/decomposition_mod.f90:64-67).  The periodic
    shallow-water model on the tile set == the serial oracle on the undivided domain, bit for bit;
    plus the halo known answer (every halo cell = the periodic global field at the wrapped position)."""
    os.environ["DL_ESM_ALIGNMENT"] = str(alignment)
    gmask = island_mask(n_x, n_y) if masked else None
    osw = O.ShallowWater(n_x, n_y, alignment=1,
                         tmask=np.asfortranarray(np.pad(gmask, 1, mode="wrap")) if masked else None)
    oi = osw.fb["t"].internal
    core = (slice(oi.xstart - 1, oi.xstop), slice(oi.ystart - 1, oi.ystop))
    init = {k: np.asfortranarray(osw.f[k][core]) for k in ("u", "v", "p")}
    for _ in range(nsteps):
        osw.step()
    ts = fd.tileset(fd.GO_ARAKAWA_C, BCS_PER, fd.GO_OFFSET_SW, n_x, n_y, ndomains=tiles_per_rank * WORLD)
    ok(ts.ntiles == tiles_per_rank and ts.tile_ids[0] == RANK * tiles_per_rank + 1, "tiles owned by this rank")

    def local(k, G, halo=0, fill=0.0):
        g = ts.grids[k]
        sb, li = g.subdomain, g.subdomain.internal
        a = np.full((g.nx, g.ny), fill, order="F")
        ji = np.arange(li.xstart - halo, li.xstop + halo + 1)
        jj = np.arange(li.ystart - halo, li.ystop + halo + 1)
        gi = (sb.glob.xstart + (ji - li.xstart) - 1) % n_x
        gj = (sb.glob.ystart + (jj - li.ystart) - 1) % n_y
        a[np.ix_(ji - 1, jj - 1)] = G[np.ix_(gi, gj)]
        return a

    for k, g in enumerate(ts.grids):
        tm = None
        if masked:
            sb = g.subdomain
            mi = (sb.glob.xstart - 2 + np.arange(sb.glob.nx)) % n_x
            mj = (sb.glob.ystart - 2 + np.arange(sb.glob.ny)) % n_y
            tm = np.asfortranarray(gmask[np.ix_(mi, mj)])
        fd.grid_init(g, 1.0e5, 1.0e5, tm)
    # halo known answer on two fields at once
    G = np.asfortranarray(np.random.default_rng(n_x * 13 + n_y).uniform(-1, 1, size=(n_x, n_y)))
    fa = [fd.r2d_field(g, fd.GO_T_POINTS) for g in ts.grids]
    fb = [fd.r2d_field(g, fd.GO_V_POINTS) for g in ts.grids]
    for k in range(ts.ntiles):
        fa[k].set_data(local(k, G, fill=-7.0))
        fb[k].set_data(local(k, -G, fill=-7.0))
    ts.halo_exchange([fa, fb])
    for k in range(ts.ntiles):
        li = ts.grids[k].subdomain.internal
        sl = (slice(li.xstart - 2, li.xstop + 1), slice(li.ystart - 2, li.ystop + 1))
        ok(np.array_equal(fa[k].get_data()[sl], local(k, G, halo=1)[sl]), "tile-set halo exchange, field 1 tile %d" % k)
        ok(np.array_equal(fb[k].get_data()[sl], local(k, -G, halo=1)[sl]), "tile-set halo exchange, field 2 tile %d" % k)
    out = ts.gather_inner_data(fa)
    if RANK == 0:
        ok(np.array_equal(out, G), "tile-set gather_inner_data")
    for f in fa + fb:
        f.free()
    # the model
    swt = fd.shallow_water_tiles(ts, use_tmask=masked)
    swt.set_state([tuple(local(k, init[n]) for n in ("u", "v", "p")) for k in range(ts.ntiles)])
    swt.step(nsteps)
    fd.dev_sync()
    for name in ("u", "v", "p", "uold", "vold", "pold"):
        out = swt.gather(name)
        if RANK == 0:
            want = osw.f[name][core]
            ok(np.array_equal(out, want), "tile-set shallow water %dx%d, %d tiles per rank, field %s: %d cells differ"
               % (n_x, n_y, tiles_per_rank, name, int((out != want).sum())))
    for a, b in zip(swt.checksums(), osw.checksums()):
        ok(abs(a - b) <= 1e-12 * abs(b), "tile-set checksum %r vs oracle %r" % (a, b))
    swt.free()
    ts.free()
    os.environ.pop("DL_ESM_ALIGNMENT", None)
    fd.gocean_initialise(RANK + 1, WORLD)


def main():
    bring_up()
    case_example_model()
    case_gsum()
    for shape in ((10, 4), (4, 10), (10, 10), (37, 23)):
        case_halos(*shape)
    case_halos(10, 10, alignment=8)
    case_halos(64, 48, alignment=64)
    case_reduction(10)
    # the same exchanges through pack -> ncclSend/ncclRecv -> unpack (the default above is the direct
    # path: patches stored straight into the neighbours' receive slots over NVLink + flag hand-shake)
    fd.set_exchange_mode(fd.EXCHANGE_NCCL)
    for shape in ((10, 4), (4, 10), (10, 10), (37, 23)):
        case_halos(*shape)
    case_halos(64, 48, alignment=64)
    case_shallow_water(130, 77, 4, 64, fd.SW_UNFUSED)
    case_shallow_water(300, 523, 8, 64, fd.SW_FUSED, two_step=True)
    fd.set_exchange_mode(fd.EXCHANGE_DIRECT)
    for mode in (fd.SW_UNFUSED, fd.SW_FUSED):
        case_shallow_water(64, 96, 5, 2, mode)
        case_shallow_water(130, 77, 4, 64, mode)
    case_shallow_water(1024, 1024, 3, 64, fd.SW_FUSED)
    case_gather()
    for shape in ((10, 10), (37, 23), (64, 48)):
        case_halos_depth2(*shape)
    case_nemolite2d(40, 36, 5, None)
    case_nemolite2d(130, 77, 4, 64)
    case_nemolite2d(257, 300, 3, 8)
    for shape in ((64, 96, 5, 2), (130, 77, 4, 64), (300, 523, 7, 64), (1024, 1024, 6, 64)):
        case_shallow_water(*shape, fd.SW_FUSED, halo_mode=fd.HALO_PUSH)
    for shape in ((64, 96, 6, 2), (130, 77, 7, 64), (300, 523, 8, 64), (1024, 1024, 9, 64)):
        case_shallow_water(*shape, fd.SW_FUSED, two_step=True)
        case_shallow_water(*shape, fd.SW_FUSED, halo_mode=fd.HALO_PUSH, two_step=True)
    # tile widths that differ by one across ranks with DL_ESM_ALIGNMENT=2: the neighbours' pitches
    # differ (x split, 4+ ranks), so every push box uses the destination's pitch
    case_shallow_water(131, 77, 6, 2, fd.SW_FUSED, halo_mode=fd.HALO_PUSH)
    case_shallow_water(131, 77, 7, 2, fd.SW_FUSED, two_step=True)
    case_shallow_water(131, 77, 7, 2, fd.SW_FUSED, halo_mode=fd.HALO_PUSH, two_step=True)
    # tmask with land points across ranks (config 3 pattern): one-step sweep, then the two-step sweep
    # whose mask ring two cells deep comes from one depth-2 exchange of the mask
    for shape in ((130, 77, 6, 64), (300, 523, 7, 64)):
        case_shallow_water(*shape, fd.SW_FUSED, masked=True)
        case_shallow_water(*shape, fd.SW_FUSED, halo_mode=fd.HALO_PUSH, masked=True)
        case_shallow_water(*shape, fd.SW_FUSED, two_step=True, masked=True)
        case_shallow_water(*shape, fd.SW_FUSED, halo_mode=fd.HALO_PUSH, two_step=True, masked=True)
    # masked one-step push into neighbours whose pitch differs (131 columns, DL_ESM_ALIGNMENT=2)
    case_shallow_water(131, 77, 6, 2, fd.SW_FUSED, halo_mode=fd.HALO_PUSH, masked=True)
    case_shallow_water(131, 77, 7, 2, fd.SW_FUSED, halo_mode=fd.HALO_PUSH, two_step=True, masked=True)
    # contracted arithmetic across ranks: within 1e-12 of the serial oracle
    for shape in ((300, 523, 8, 64), (1024, 1024, 9, 64)):
        case_shallow_water(*shape, fd.SW_FUSED, halo_mode=fd.HALO_PUSH, two_step=True, arith=fd.ARITH_CONTRACTED)
        case_shallow_water(*shape, fd.SW_FUSED, halo_mode=fd.HALO_PUSH, arith=fd.ARITH_CONTRACTED)
    # several sub-domains per rank (extension F3): 2 and 3 tiles per rank, with and without land
    case_tileset(64, 96, 2, 5, 2)
    case_tileset(130, 77, 3, 4, 64)
    case_tileset(300, 523, 2, 6, 64, masked=True)
    case_tileset(1024, 1024, 2, 5, 64)
    dist.barrier()
    fd.gocean_finalise()
    if RANK == 0:
        print("DIST OK %d checks on %d ranks" % (NCHECK, WORLD), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
