"""CPU tests: the product's host-side bookkeeping (C++ behind the C ABI, no GPU needed)
against the oracle, bit-exact, over the reference's own configurations and many more."""
import itertools
import os

import numpy as np
import pytest

import dl_esm_inf_b200 as fd
from oracle import oracle as O

BCS_EXT = (fd.GO_BC_EXTERNAL, fd.GO_BC_EXTERNAL, fd.GO_BC_NONE)
BCS_PER = (fd.GO_BC_PERIODIC, fd.GO_BC_PERIODIC, fd.GO_BC_NONE)


def region_eq(a, b):
    return a.astuple() == b.astuple()


def box_eq(a, b):
    return a.box() == b.box()


DECOMP_CASES = [
    # (domainx, domainy, nranks, kwargs)
    (4, 10, 1, {}), (4, 10, 2, {}), (10, 4, 2, {}), (4, 10, 4, {}), (4, 10, 6, {}),
    (10, 10, 4, {}), (10, 10, 6, {}), (10, 10, 3, {}), (10, 10, 7, {}), (10, 10, 8, {}),
    (16384, 16384, 2, {}), (16384, 16384, 4, {}), (16384, 16384, 8, {}),
    (16384, 32768, 2, {}), (32768, 32768, 4, {}), (32768, 65536, 8, {}),
    (33, 17, 5, {}), (17, 33, 12, {}), (100, 7, 9, {}), (1, 1, 1, {}), (5, 3, 16, {}),
    (64, 64, 1, dict(ndomains=4)), (64, 48, 1, dict(ndomainx=3, ndomainy=2)),
    (64, 48, 6, dict(ndomainx=2, ndomainy=3)), (50, 50, 4, dict(halo_width=2)),
    (50, 50, 1, dict(halo_width=0)), (1024, 1024, 1, {}),
]


@pytest.mark.parametrize("dx_,dy_,nranks,kw", DECOMP_CASES)
def test_go_decompose_matches_oracle(dx_, dy_, nranks, kw):
    O.reset_saved_state()
    fd.gocean_initialise(1, nranks)
    want = O.go_decompose(dx_, dy_, nranks=nranks, **kw)
    got = fd.go_decompose(dx_, dy_, **kw)
    assert (got.global_nx, got.global_ny, got.nx, got.ny, got.ndomains, got.max_width,
            got.max_height, got.domperrank) == (
        want.global_nx, want.global_ny, want.nx, want.ny, want.ndomains, want.max_width,
        want.max_height, want.domperrank)
    subs = got.subdomains
    for k in range(want.ndomains):
        assert region_eq(subs[k].glob, want.subdomains[k].glob), k
        assert region_eq(subs[k].internal, want.subdomains[k].internal), k
    for r in range(1, nranks + 1):
        for s in range(1, want.domperrank + 1):
            assert got.proc_subdomains(s, r) == want.proc_subdomains[(r - 1) * want.domperrank + (s - 1)]


def test_go_decompose_saved_halo_width():
    """halo_width sticks between calls (SAVEd local, src/parallel_mod.f90:103-105)."""
    O.reset_saved_state()
    fd.gocean_initialise(1, 1)
    for kw in (dict(halo_width=3), {}):
        want = O.go_decompose(20, 20, nranks=1, **kw)
        got = fd.go_decompose(20, 20, **kw)
        assert got.subdomains[0].internal.xstart == want.subdomains[0].internal.xstart == 4


def test_go_decompose_errors():
    fd.gocean_initialise(1, 2)
    with pytest.raises(fd.GoceanStop) as e:
        fd.go_decompose(10, 10, ndomainx=2)
    assert e.value.code == O.ERR_ARGS
    with pytest.raises(O.OracleStop) as eo:
        O.go_decompose(10, 10, ndomainx=2, nranks=2)
    assert eo.value.code == e.value.code
    with pytest.raises(fd.GoceanStop) as e:
        fd.go_decompose(10, 10, halo_width=0)
    with pytest.raises(O.OracleStop) as eo:
        O.go_decompose(10, 10, halo_width=0, nranks=2)
    assert e.value.code == eo.value.code == O.ERR_HALO_WIDTH


@pytest.mark.parametrize("dx_,dy_,nranks", [(4, 10, 2), (10, 4, 2), (10, 10, 4), (10, 10, 6),
                                             (4, 10, 4), (4, 10, 6), (64, 48, 12), (33, 17, 5),
                                             (16384, 16384, 8), (100, 7, 9), (5, 3, 15), (9, 9, 9)])
def test_map_comms_matches_oracle_on_every_rank(dx_, dy_, nranks):
    O.reset_saved_state()
    fd.gocean_initialise(1, nranks)
    want_d = O.go_decompose(dx_, dy_, nranks=nranks)
    got_d = fd.go_decompose(dx_, dy_)
    names = [n for n, _ in fd.Comms._fields_ if n not in ("nsend", "nrecv")]
    for rank in range(1, nranks + 1):
        want = O.map_comms(want_d, rank, nranks)
        got = fd.map_comms(got_d, rank, nranks)
        assert (got.nsend, got.nrecv) == (want.nsend, want.nrecv)
        for n in names:
            cnt = want.nsend if n in ("dirsend", "destination", "isrcsend", "jsrcsend", "idessend",
                                      "jdessend", "nxsend", "nysend", "nsendp") else want.nrecv
            assert list(getattr(got, n))[:cnt] == list(getattr(want, n))[:cnt], (rank, n)
            # unused slots keep the reference's -999 initialisation
            assert all(v == -999 for v in list(getattr(got, n))[cnt:]), (rank, n)
        for ia, ja in itertools.product((0, 1, dx_ // 2 + 1, dx_, dx_ + 1), (0, 1, dy_ // 2 + 1, dy_, dy_ + 1)):
            assert fd.iprocmap(got_d, ia, ja, nranks) == O.lib().orc_iprocmap(want_d, nranks, ia, ja)


def test_map_comms_maxcomm_and_depth_errors():
    fd.gocean_initialise(1, 4)
    d = fd.go_decompose(10, 10)
    with pytest.raises(fd.GoceanStop) as e:
        fd.map_comms(d, 1, 4, halo_depths=(2, 2))
    assert e.value.code == O.ERR_HALO_DEPTH
    with pytest.raises(O.OracleStop) as eo:
        O.map_comms(O.go_decompose(10, 10, nranks=4), 1, 4, halo_depths=(2, 2))
    assert eo.value.code == O.ERR_HALO_DEPTH


@pytest.mark.parametrize("env,ok", [(None, 1), ("1", 1), ("64", 64), ("8", 8), ("128", 128), (" 16", 16),
                                    ("999", 999), ("1000", O.ERR_ALIGN_DIGITS), ("0", O.ERR_ALIGN_VALUE),
                                    ("-4", O.ERR_ALIGN_VALUE), ("abc", O.ERR_ALIGN_VALUE),
                                    ("4x", O.ERR_ALIGN_VALUE)])
def test_alignment_env_matches_oracle(env, ok, monkeypatch):
    if env is None:
        monkeypatch.delenv("DL_ESM_ALIGNMENT", raising=False)
    else:
        monkeypatch.setenv("DL_ESM_ALIGNMENT", env)
    fd.gocean_initialise(1, 1)
    g = fd.grid_type(fd.GO_ARAKAWA_C, BCS_EXT, fd.GO_OFFSET_NE)
    g.decompose(5, 5)
    if ok > 0:
        assert O.parse_alignment(env) == ok
        fd.grid_init(g, 1.0, 1.0)
        gd = O.GridDims()
        sub = O.go_decompose(5, 5).subdomains[0]
        O.lib().orc_grid_dims_compute(sub, ok, gd)
        assert (g.nx, g.ny, g.info.padding, g.info.alignment) == (gd.nx, gd.ny, gd.padding, ok)
        assert g.nx % ok == 0 and g.nx > sub.glob.nx
    else:
        with pytest.raises(O.OracleStop) as eo:
            O.parse_alignment(env)
        with pytest.raises(fd.GoceanStop) as e:
            fd.grid_init(g, 1.0, 1.0)
        assert e.value.code == eo.value.code == ok


GRID_CASES = [
    # (nx, ny, nranks, offset, bcs, alignment, with_tmask, dx, dy)
    (4, 10, 1, fd.GO_OFFSET_NE, BCS_EXT, 1, True, 1.0, 1.0),
    (4, 10, 2, fd.GO_OFFSET_NE, BCS_EXT, 1, True, 1.0, 1.0),
    (10, 10, 6, fd.GO_OFFSET_NE, BCS_EXT, 1, True, 1.0, 1.0),
    (10, 10, 4, fd.GO_OFFSET_NE, BCS_EXT, 1, False, 1.0, 1.0),
    (5, 5, 1, fd.GO_OFFSET_NE, BCS_EXT, 64, True, 1.0, 1.0),
    (37, 23, 6, fd.GO_OFFSET_NE, BCS_EXT, 8, True, 0.1, 0.3),
    (64, 48, 1, fd.GO_OFFSET_SW, BCS_PER, 1, False, 1.0e5, 1.0e5),
    (64, 48, 1, fd.GO_OFFSET_SW, BCS_PER, 64, True, 1000.0, 2000.0),
    (30, 30, 1, fd.GO_OFFSET_SW, (fd.GO_BC_EXTERNAL, fd.GO_BC_PERIODIC, fd.GO_BC_NONE), 1, True, 1.0, 1.0),
    (1024, 1024, 1, fd.GO_OFFSET_SW, BCS_PER, 1, False, 1.0e5, 1.0e5),
]


def _random_tmask(sub, seed):
    rng = np.random.default_rng(seed)
    return np.asfortranarray(rng.integers(-1, 2, size=(sub.glob.nx, sub.glob.ny)).astype(np.int32))


@pytest.mark.parametrize("nx_,ny_,nranks,offset,bcs,align,with_tmask,dx,dy", GRID_CASES)
def test_grid_init_and_field_bounds_match_oracle(nx_, ny_, nranks, offset, bcs, align, with_tmask,
                                                 dx, dy, monkeypatch):
    monkeypatch.setenv("DL_ESM_ALIGNMENT", str(align))
    for rank in range(1, nranks + 1):
        O.reset_saved_state()
        fd.gocean_initialise(rank, nranks)
        og = O.Grid(O.GO_ARAKAWA_C, bcs, offset)
        og.decompose(nx_, ny_, rank=rank, nranks=nranks)
        tm = _random_tmask(og.subdomain, 100 * rank + nx_) if with_tmask else None
        og.init(dx, dy, tmask=tm, alignment=align)
        g = fd.grid_type(fd.GO_ARAKAWA_C, bcs, offset)
        g.decompose(nx_, ny_)
        fd.grid_init(g, dx, dy, tmask=tm)
        assert (g.nx, g.ny, g.global_nx, g.global_ny) == (og.nx, og.ny, og.global_nx, og.global_ny)
        assert region_eq(g.subdomain.glob, og.subdomain.glob)
        assert region_eq(g.subdomain.internal, og.subdomain.internal)
        assert (g.dx, g.dy) == (og.dx, og.dy)
        # tmask: everywhere when a mask was supplied; inside the initialised box otherwise
        got_t = g.get_tmask()
        if with_tmask:
            assert np.array_equal(got_t, og.tmask)
        else:
            defined = og.tmask != O.TMASK_UNSET
            assert defined.any() and np.array_equal(got_t[defined], og.tmask[defined])
        # metrics: bit-exact, including the running-sum coordinates
        if nx_ <= 64:
            for name in O.METRIC_NAMES:
                assert np.array_equal(getattr(g, name), og.metrics[name]), name
        else:
            for name in ("xt", "yt", "area_t"):
                assert np.array_equal(getattr(g, name), og.metrics[name]), name
        # comm tables installed by grid_init (only with more than one rank)
        if nranks > 1:
            assert list(g.comms.nxsend)[:g.comms.nsend] == list(og.comms.nxsend)[:og.comms.nsend]
        # field bounds for every point type (device storage is not touched: bounds only)
        for pts in (fd.GO_U_POINTS, fd.GO_V_POINTS, fd.GO_T_POINTS, fd.GO_F_POINTS, fd.GO_ALL_POINTS):
            try:
                want = O.field_bounds(og, pts)
                want_err = None
            except O.OracleStop as e:
                want, want_err = None, e.code
            fb = _product_bounds(g, pts)
            if want_err is not None:
                assert fb == want_err, pts
                continue
            assert not isinstance(fb, int), (pts, fb)
            assert fb.defined_on == want.defined_on
            assert region_eq(fb.internal, want.internal) and region_eq(fb.whole, want.whole)
            assert fb.num_halos == want.num_halos
            for k in range(max(want.num_halos, 0)):
                assert box_eq(fb.halo[k].source, want.halo[k].source)
                assert box_eq(fb.halo[k].dest, want.halo[k].dest)


def _product_bounds(g, pts):
    """set_field_bounds through the C ABI (host only); gocean_stop branches give their code"""
    try:
        return g.field_bounds(pts)
    except fd.GoceanStop as e:
        return e.code


@pytest.mark.parametrize("dx_,dy_,nranks", [(64, 48, 1), (64, 48, 2), (40, 90, 3), (64, 48, 4), (96, 50, 6), (128, 96, 8),
                                             (99, 99, 9), (64, 300, 12), (256, 256, 16)])
@pytest.mark.parametrize("depth", [1, 2])
def test_direction_tags_address_one_slot_per_message(dx_, dy_, nranks, depth):
    """What the direct exchange and the packed push strips of the two-step sweep rely on: with
    periodic tables of any depth, every send entry of rank r meets EXACTLY ONE receive entry on its destination
    with source r, the same direction tag and the same size -- so the tag names a slot both ends agree on -- and
    no two messages into one rank carry the same tag (a slot is never shared).  Narrow boxes (at most `depth`
    columns wide: what leaves as a packed strip) sit on the tile's two x-edge column pairs."""
    fd.gocean_initialise(1, nranks)
    fd.set_extensions(fd.EXT_PERIODIC_DISTMEM | fd.EXT_HALO_DEPTH)
    try:
        d = fd.go_decompose(dx_, dy_, halo_width=depth)
        tabs = {r: fd.map_comms(d, r, nranks, pbc=(True, True), halo_depths=(depth, depth)) for r in range(1, nranks + 1)}
        for r, t in tabs.items():
            sub = d.subdomains[r - 1].internal if hasattr(d, "subdomains") else None
            tags_in = [t.dirrecv[q] for q in range(t.nrecv)]
            assert len(set(tags_in)) == len(tags_in), (r, tags_in)          # one message per tag into a rank
            assert all(1 <= g <= 8 for g in tags_in)
            for s in range(t.nsend):
                dest = t.destination[s] + 1 if nranks > 1 else 1     # the tables hold 0-based ranks
                td = tabs[dest]
                hits = [q for q in range(td.nrecv)
                        if (td.source[q] == r - 1 or nranks == 1) and td.dirrecv[q] == t.dirsend[s]]
                assert len(hits) == 1, (r, s, dest, hits)
                q = hits[0]
                assert (td.nxrecv[q], td.nyrecv[q]) == (t.nxsend[s], t.nysend[s])
                if sub is not None and t.nxsend[s] <= depth and sub.nx >= 2 * depth:
                    i0 = t.isrcsend[s]
                    assert i0 == sub.xstart or i0 + t.nxsend[s] - 1 == sub.xstop, (r, s, i0)
    finally:
        fd.set_extensions(0)
        fd.gocean_initialise(1, 1)


def test_go_decompose_and_map_comms_match_oracle_on_random_cases():
    """property test: the product's independent C++ bookkeeping (cut vectors + owner lookup, tensor-product
    message tables) equals the oracle's loop-for-loop restatement of src/parallel_mod.f90:95-332 and
    src/parallel_comms_mod.f90:178-1172 on random domains, rank counts and halo widths -- or both stop with
    the same gocean_stop code"""
    from hypothesis import given, settings, strategies as st, HealthCheck

    names = [n for n, _ in fd.Comms._fields_ if n not in ("nsend", "nrecv")]
    send_names = ("dirsend", "destination", "isrcsend", "jsrcsend", "idessend", "jdessend", "nxsend", "nysend", "nsendp")

    @settings(max_examples=120, deadline=None, suppress_health_check=list(HealthCheck))
    @given(dx_=st.integers(1, 400), dy_=st.integers(1, 400), nranks=st.integers(1, 16), hw=st.integers(1, 3))
    def check(dx_, dy_, nranks, hw):
        O.reset_saved_state()
        fd.gocean_initialise(1, nranks)
        try:
            want_d = O.go_decompose(dx_, dy_, nranks=nranks, halo_width=hw)
            want_err = None
        except O.OracleStop as e:
            want_d, want_err = None, e.code
        try:
            got_d = fd.go_decompose(dx_, dy_, halo_width=hw)
            got_err = None
        except fd.GoceanStop as e:
            got_d, got_err = None, e.code
        assert got_err == want_err, (dx_, dy_, nranks, hw)
        if want_err is not None:
            return
        assert (got_d.nx, got_d.ny, got_d.ndomains, got_d.max_width, got_d.max_height) == (
            want_d.nx, want_d.ny, want_d.ndomains, want_d.max_width, want_d.max_height)
        subs = got_d.subdomains
        for k in range(want_d.ndomains):
            assert region_eq(subs[k].glob, want_d.subdomains[k].glob), k
            assert region_eq(subs[k].internal, want_d.subdomains[k].internal), k
        # message tables: only where every rank owns at least one point (a domain smaller than the rank grid
        # leaves EMPTY sub-domains, for which the reference's border walk has no defined answer)
        if any(sd.internal.nx < 1 or sd.internal.ny < 1 for sd in want_d.subdomains[:want_d.ndomains]):
            return
        for rank in range(1, nranks + 1):
            try:
                want = O.map_comms(want_d, rank, nranks)
                werr = None
            except O.OracleStop as e:
                want, werr = None, e.code
            try:
                got = fd.map_comms(got_d, rank, nranks)
                gerr = None
            except fd.GoceanStop as e:
                got, gerr = None, e.code
            assert gerr == werr, (dx_, dy_, nranks, hw, rank)
            if werr is not None:
                continue
            assert (got.nsend, got.nrecv) == (want.nsend, want.nrecv)
            for n in names:
                cnt = want.nsend if n in send_names else want.nrecv
                w = list(getattr(want, n))[:cnt]
                if hw > 1 and n in ("idessend", "jdessend", "idesrecv", "jdesrecv"):
                    # INTEGRATION.md 4, difference 3: the reference addresses a low-side halo as 1..depth
                    # (src/parallel_comms_mod.f90:334,524 -- written for halo width == depth); the product anchors
                    # it at the internal region's edge, which is what a stencil reads
                    w = [hw if v == 1 else v for v in w]
                assert list(getattr(got, n))[:cnt] == w, (dx_, dy_, nranks, hw, rank, n)

    check()
    O.reset_saved_state()
    fd.gocean_initialise(1, 1)


def test_field_bounds_match_oracle_on_random_grids(monkeypatch):
    """property test: grid sizing (DL_ESM_ALIGNMENT padding, src/grid_mod.f90:368-385) and set_field_bounds for
    every point type, offset and boundary-condition mix (src/field_mod.f90:563-1122, incl. the gocean_stop
    branches and the periodic halo lists) equal the oracle on random domains"""
    from hypothesis import given, settings, strategies as st, HealthCheck

    bc = st.sampled_from([fd.GO_BC_EXTERNAL, fd.GO_BC_PERIODIC])

    @settings(max_examples=80, deadline=None, suppress_health_check=list(HealthCheck))
    @given(nx_=st.integers(2, 150), ny_=st.integers(2, 150), offset=st.sampled_from([fd.GO_OFFSET_NE, fd.GO_OFFSET_SW]),
           bcx=bc, bcy=bc, align=st.sampled_from([1, 2, 8, 64]))
    def check(nx_, ny_, offset, bcx, bcy, align):
        monkeypatch.setenv("DL_ESM_ALIGNMENT", str(align))
        O.reset_saved_state()
        fd.gocean_initialise(1, 1)
        bcs = (bcx, bcy, fd.GO_BC_NONE)
        og = O.Grid(O.GO_ARAKAWA_C, bcs, offset)
        og.decompose(nx_, ny_, rank=1, nranks=1)
        g = fd.grid_type(fd.GO_ARAKAWA_C, bcs, offset)
        g.decompose(nx_, ny_)
        try:
            og.init(1.0, 2.0, tmask=None, alignment=align)
            want_err = None
        except O.OracleStop as e:
            want_err = e.code
        try:
            fd.grid_init(g, 1.0, 2.0)
            got_err = None
        except fd.GoceanStop as e:
            got_err = e.code
        assert got_err == want_err, (nx_, ny_, offset, bcx, bcy, align)
        if want_err is not None:
            return
        assert (g.nx, g.ny) == (og.nx, og.ny) and g.nx % align == 0
        for pts in (fd.GO_U_POINTS, fd.GO_V_POINTS, fd.GO_T_POINTS, fd.GO_F_POINTS, fd.GO_ALL_POINTS):
            try:
                want, werr = O.field_bounds(og, pts), None
            except O.OracleStop as e:
                want, werr = None, e.code
            fb = _product_bounds(g, pts)
            if werr is not None:
                assert fb == werr, (nx_, ny_, offset, bcx, bcy, pts)
                continue
            assert not isinstance(fb, int), (nx_, ny_, offset, bcx, bcy, pts, fb)
            assert region_eq(fb.internal, want.internal) and region_eq(fb.whole, want.whole)
            assert fb.num_halos == want.num_halos
            for k in range(max(want.num_halos, 0)):
                assert box_eq(fb.halo[k].source, want.halo[k].source) and box_eq(fb.halo[k].dest, want.halo[k].dest)

    check()
    fd.gocean_initialise(1, 1)


def test_comm_plans_pair_up_across_ranks_on_random_cases():
    """property test of the NCCL-mode message plan (dlesm_comm_plan_build) without any communicator: for random
    domains, rank counts (up to 16), periodicity and depth, what rank a plans to send to rank b is exactly what b
    plans to receive from a -- the same number of messages, the same element count, segment by segment in the
    order both ends derive from the direction tags"""
    from hypothesis import given, settings, strategies as st, HealthCheck
    from dl_esm_inf_b200.api import comm_plan

    @settings(max_examples=60, deadline=None, suppress_health_check=list(HealthCheck))
    @given(dx_=st.integers(8, 300), dy_=st.integers(8, 300), nranks=st.integers(2, 16), depth=st.integers(1, 2),
           periodic=st.booleans())
    def check(dx_, dy_, nranks, depth, periodic):
        fd.gocean_initialise(1, nranks)
        fd.set_extensions((fd.EXT_PERIODIC_DISTMEM if periodic else 0) | fd.EXT_HALO_DEPTH)
        try:
            d = fd.go_decompose(dx_, dy_, halo_width=depth)
            subs = d.subdomains
            if any(subs[k].internal.nx < depth or subs[k].internal.ny < depth for k in range(nranks)):
                return                      # a tile thinner than the halo it has to send
            plans = {}
            for r in range(1, nranks + 1):
                c = fd.map_comms(d, r, nranks, pbc=(periodic, periodic), halo_depths=(depth, depth))
                plans[r] = (c, comm_plan(c, r))
            for a, (ca, pa) in plans.items():
                for k in range(pa.npeers):
                    b = pa.peer[k] + 1
                    if b == a:
                        continue
                    cb, pb = plans[b]
                    back = [q for q in range(pb.npeers) if pb.peer[q] + 1 == a]
                    assert len(back) == 1, (a, b)
                    q = back[0]
                    assert pa.nsend[k] == pb.nrecv[q] and pa.send_n[k] == pb.recv_n[q], (a, b)
                    for m in range(pa.nsend[k]):
                        s, rcv = pa.send_slot[k][m], pb.recv_slot[q][m]
                        assert ca.nxsend[s] * ca.nysend[s] == cb.nxrecv[rcv] * cb.nyrecv[rcv], (a, b, m)
                        assert ca.dirsend[s] == cb.dirrecv[rcv], (a, b, m)
        finally:
            fd.set_extensions(0)

    check()
    fd.gocean_initialise(1, 1)


def test_periodic_tables_reproduce_the_global_field_on_random_cases():
    """property test: an exchange SIMULATED in one process from the product's map_comms tables of every rank
    (send boxes copied into the destination rank's array at the table's (ides, jdes)) leaves in every halo cell
    within `depth` of the internal region -- corners included -- the value of the periodic global field at the
    wrapped position.  Random domains, 1..16 ranks, depth 1..3: the index arithmetic of the wrap-around
    (DLESM_EXT_PERIODIC_DISTMEM) and deep-halo (DLESM_EXT_HALO_DEPTH) tables, at rank counts no test box has."""
    from hypothesis import given, settings, strategies as st, HealthCheck

    @settings(max_examples=60, deadline=None, suppress_health_check=list(HealthCheck))
    @given(dx_=st.integers(6, 120), dy_=st.integers(6, 120), nranks=st.integers(1, 16), depth=st.integers(1, 3),
           seed=st.integers(0, 10 ** 6))
    def check(dx_, dy_, nranks, depth, seed):
        fd.gocean_initialise(1, nranks)
        fd.set_extensions(fd.EXT_PERIODIC_DISTMEM | fd.EXT_HALO_DEPTH)
        try:
            d = fd.go_decompose(dx_, dy_, halo_width=depth)
            subs = [d.subdomains[k] for k in range(nranks)]
            if any(sd.internal.nx < depth or sd.internal.ny < depth for sd in subs):
                return
            G = np.random.default_rng(seed).uniform(-1, 1, size=(dx_, dy_))
            arrs = []
            for sd in subs:
                li = sd.internal
                a = np.full((sd.glob.nx, sd.glob.ny), -7.0)
                gi = sd.glob.xstart + np.arange(li.nx) - 1
                gj = sd.glob.ystart + np.arange(li.ny) - 1
                a[li.xstart - 1:li.xstop, li.ystart - 1:li.ystop] = G[np.ix_(gi, gj)]
                arrs.append(a)
            tabs = [fd.map_comms(d, r, nranks, pbc=(True, True), halo_depths=(depth, depth)) for r in range(1, nranks + 1)]
            # all sources are internal cells, all destinations halo cells: the order of the copies does not matter
            for r, t in enumerate(tabs):
                for s in range(t.nsend):
                    dest = t.destination[s] if nranks > 1 else 0
                    i0, j0, nx, ny = t.isrcsend[s] - 1, t.jsrcsend[s] - 1, t.nxsend[s], t.nysend[s]
                    di, dj = t.idessend[s] - 1, t.jdessend[s] - 1
                    arrs[dest][di:di + nx, dj:dj + ny] = arrs[r][i0:i0 + nx, j0:j0 + ny]
            for sd, a in zip(subs, arrs):
                li = sd.internal
                ji = np.arange(li.xstart - depth, li.xstop + depth + 1)
                jj = np.arange(li.ystart - depth, li.ystop + depth + 1)
                wi = (sd.glob.xstart + (ji - li.xstart) - 1) % dx_
                wj = (sd.glob.ystart + (jj - li.ystart) - 1) % dy_
                assert np.array_equal(a[np.ix_(ji - 1, jj - 1)], G[np.ix_(wi, wj)]), (dx_, dy_, nranks, depth)
        finally:
            fd.set_extensions(0)

    check()
    fd.gocean_initialise(1, 1)


def test_tileset_plans_reproduce_the_global_field_on_random_cases():
    """property test of the several-sub-domains-per-rank plan (dlesm_tileset_plan, extension F3): the plans of
    EVERY rank, executed in one process (same-rank copies; per-peer buffers handed from the sender's plan to the
    receiver's), fill every halo cell inside the -- periodic: wrapped -- domain with the global field's value.
    Random domains, 1..6 ranks x 1..4 tiles, depth 1..2."""
    from hypothesis import given, settings, strategies as st, HealthCheck
    from dl_esm_inf_b200.api import tileset_plan

    @settings(max_examples=50, deadline=None, suppress_health_check=list(HealthCheck))
    @given(nx=st.integers(12, 96), ny=st.integers(12, 96), world=st.integers(1, 6), tpr=st.integers(1, 4),
           periodic=st.booleans(), depth=st.integers(1, 2), seed=st.integers(0, 10 ** 6))
    def check(nx, ny, world, tpr, periodic, depth, seed):
        fd.gocean_initialise(1, world)
        d = fd.go_decompose(nx, ny, ndomains=tpr * world, halo_width=depth)
        subs = d.subdomains
        if any(subs[k].internal.nx < depth or subs[k].internal.ny < depth for k in range(tpr * world)):
            return
        G = np.random.default_rng(seed).uniform(-1, 1, size=(nx, ny))
        arrs, plans, sendbufs, recvbufs = {}, {}, {}, {}
        for r in range(world):
            owned = [d.proc_subdomains(s, r + 1) for s in range(1, tpr + 1)]
            local = []
            for tid in owned:
                sb = subs[tid - 1]
                li = sb.internal
                a = np.full((sb.glob.nx, sb.glob.ny), -7.0, order="F")
                gi = sb.glob.xstart + np.arange(li.nx) - 1
                gj = sb.glob.ystart + np.arange(li.ny) - 1
                a[li.xstart - 1:li.xstop, li.ystart - 1:li.ystop] = G[np.ix_(gi, gj)]
                local.append(a)
            arrs[r] = (owned, local)
            rows, (stot, rtot) = tileset_plan(d, r + 1, world, pbc=(periodic, periodic), depth=depth)
            plans[r] = rows.tolist()
            sendbufs[r], recvbufs[r] = np.zeros(stot), np.zeros(rtot)
        region_s, region_r = {}, {}
        for r in range(world):
            before = [a.copy() for a in arrs[r][1]]
            for kind, peer, st_, dt_, si0, sj0, di0, dj0, w, h, off, dst_id, src_id, dirn in plans[r]:
                if kind == 0:
                    assert peer == r
                    arrs[r][1][dt_][di0 - 1:di0 - 1 + w, dj0 - 1:dj0 - 1 + h] = before[st_][si0 - 1:si0 - 1 + w, sj0 - 1:sj0 - 1 + h]
                elif kind == 1:
                    sendbufs[r][off:off + w * h] = before[st_][si0 - 1:si0 - 1 + w, sj0 - 1:sj0 - 1 + h].flatten(order="F")
                    lo, n = region_s.get((r, peer), (off, 0))
                    region_s[(r, peer)] = (min(lo, off), n + w * h)
                else:
                    lo, n = region_r.get((r, peer), (off, 0))
                    region_r[(r, peer)] = (min(lo, off), n + w * h)
        assert {(a, b) for a, b in region_s} == {(b, a) for a, b in region_r}
        for (a, b), (lo, n) in region_s.items():
            rlo, rn = region_r[(b, a)]
            assert rn == n, (a, b)
            recvbufs[b][rlo:rlo + n] = sendbufs[a][lo:lo + n]
        for r in range(world):
            for kind, peer, st_, dt_, si0, sj0, di0, dj0, w, h, off, dst_id, src_id, dirn in plans[r]:
                if kind == 2:
                    arrs[r][1][dt_][di0 - 1:di0 - 1 + w, dj0 - 1:dj0 - 1 + h] = recvbufs[r][off:off + w * h].reshape((w, h), order="F")
        for r in range(world):
            for tid, a in zip(*arrs[r]):
                sb = subs[tid - 1]
                li = sb.internal
                ji = np.arange(li.xstart - depth, li.xstop + depth + 1)
                jj = np.arange(li.ystart - depth, li.ystop + depth + 1)
                gx = sb.glob.xstart + ji - li.xstart
                gy = sb.glob.ystart + jj - li.ystart
                if periodic:
                    want = G[np.ix_((gx - 1) % nx, (gy - 1) % ny)]
                else:
                    inx, iny = (gx >= 1) & (gx <= nx), (gy >= 1) & (gy <= ny)
                    want = np.full((len(ji), len(jj)), -7.0)
                    want[np.ix_(inx, iny)] = G[np.ix_(gx[inx] - 1, gy[iny] - 1)]
                assert np.array_equal(a[np.ix_(ji - 1, jj - 1)], want), (nx, ny, world, tpr, periodic, depth, tid)

    check()
    fd.gocean_initialise(1, 1)
